/* sift3d_cuda.h -- the thin C ABI between the host C library (libsift3D.so, the
 * drop-in for the reference's public API) and the CUDA layer (libsift3d_cuda.so,
 * hand-written sm_100a kernels). Plain C types only: no CUDA, C++ or torch types
 * cross this boundary. Every function returns 0 on success and -1 on failure
 * (the reference's SIFT3D_SUCCESS / SIFT3D_FAILURE, imtypes.h:21-25) and reports
 * the reason on stderr, like the reference's SIFT3D_ERR (immacros.h:27-32).
 *
 * What each entry point replaces in the reference (paths under /root/reference):
 *
 *   sift3d_cuda_plan            resize_SIFT3D sift3d/sift.c:427-475, resize_Pyramid
 *                               sift3d/imutil.c:1464-1554 (allocation of all levels)
 *   sift3d_cuda_detect          set_im_SIFT3D sift.c:629-659 (im_scale imutil.c:699),
 *                               build_gpyr sift.c:662-711 (apply_Sep_FIR_filter
 *                               imutil.c:1127-1206, im_downsample_2x imutil.c:591),
 *                               build_dog sift.c:713-732, detect_extrema sift.c:735-871,
 *                               assign_orientations sift.c:1109-1167 (per-candidate
 *                               part; the in-order compaction stays in host C)
 *   sift3d_cuda_describe        do_extract_descriptors sift.c:1561-1596 /
 *                               extract_descrip sift.c:1442-1536
 *   sift3d_cuda_blur            apply_Sep_FIR_filter imutil.c:1127-1206 (one blur;
 *                               exposed for parity tests)
 *
 * There is NO CPU fallback behind this interface: if no CUDA device is usable
 * sift3d_cuda_ctx_create() returns NULL and every caller fails loudly.
 */
#ifndef SIFT3D_CUDA_H
#define SIFT3D_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SIFT3D_CUDA_EXPORT __attribute__((visibility("default")))

#define SIFT3D_CUDA_MAX_FILTERS 16  /* incremental filters per octave (num_kp_levels + 2) */
#define SIFT3D_CUDA_MAX_OCTAVES 24
#define SIFT3D_CUDA_MAX_TAPS 129    /* 2 * half_width + 1 */
#define SIFT3D_CUDA_DESC_NUMEL 768  /* 4*4*4 histograms x 12 icosahedron vertices */

typedef struct sift3d_cuda_ctx sift3d_cuda_ctx;

/* One separable FIR filter (imtypes_private.h:84-95). */
typedef struct {
    int width;                          /* 2*half_width + 1 */
    float taps[SIFT3D_CUDA_MAX_TAPS];   /* normalised, as init_Gauss_filter computes them */
} sift3d_cuda_filter;

/* Geometry + filters of one pyramid; filled by host C (sift3d_host.c) with the
 * reference's expressions so that every double->float conversion happens on the
 * host exactly as in the reference. */
typedef struct {
    int nx, ny, nz;                 /* octave-0 dims */
    double units[3];                /* voxel units the pyramid was (re)sized with;
                                       octave o >= 1 uses units * 2^o (imutil.c:1545-1548) */
    int num_octaves;                /* sift.c:442-454 */
    int num_kp_levels;              /* gpyr has num_kp_levels + 3 levels, s = -1 .. kp+1 */
    double scales[SIFT3D_CUDA_MAX_OCTAVES][SIFT3D_CUDA_MAX_FILTERS + 1];
                                    /* [o][i]: scale of level s = -1 + i of octave o,
                                       sigma0 * pow(2, o + s/num_kp_levels) (imutil.c:1578-1579) */
    sift3d_cuda_filter first;       /* sigma_n -> scale(-1) (imutil.c:1392-1396) */
    sift3d_cuda_filter octave[SIFT3D_CUDA_MAX_FILTERS]; /* [i]: level i-1 -> level i */
} sift3d_cuda_plan_desc;

/* A keypoint / candidate as it crosses the ABI (imtypes_private.h:159-166). */
typedef struct {
    int32_t o, s;
    double x, y, z;     /* octave-local voxel coordinates (integers, sift.c:861-863) */
    double sd;          /* level scale */
    float strength;     /* |DoG| at the candidate */
    float R[9];         /* row-major rotation matrix */
    int32_t valid;      /* 1 = survived orientation assignment */
} sift3d_cuda_keypoint;

/* Per-stage device times of the last detect/describe call, in milliseconds
 * (CUDA events on the context's stream). */
typedef struct {
    float h2d, scale, pyramid, dogmax, extrema, order, orient, d2h, total_detect;
    float desc_h2d, descriptors, desc_d2h, total_describe;
} sift3d_cuda_timings;

SIFT3D_CUDA_EXPORT int sift3d_cuda_device_count(void);

/* device < 0: use $SIFT3D_CUDA_DEVICE or 0. NULL on failure. */
SIFT3D_CUDA_EXPORT sift3d_cuda_ctx *sift3d_cuda_ctx_create(int device);
SIFT3D_CUDA_EXPORT void sift3d_cuda_ctx_destroy(sift3d_cuda_ctx *ctx);

/* Arithmetic of the Gaussian blurs (apply_Sep_FIR_filter imutil.c:1127-1206).
 *   exact != 0  every tap is a separate fp32 multiply and add in the reference's order and
 *               im_scale (imutil.c:699-713) is a true division: pyramid levels are
 *               bit-identical (0 ULP) to the reference build (x86-64 without FMA);
 *   exact == 0  (default) one fused multiply-add per tap in the fused unit-spacing kernel and
 *               im_scale folded into the first blur as a multiply by 1 / max: ~1e-7
 *               relative to the level's max-abs, inside the 1e-5 voxel tolerance of the
 *               parity contract, at half the FP32 instructions per voxel.
 * Default: $SIFT3D_CUDA_EXACT (unset or 0: fast). Takes effect at the next detect. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_set_exact(sift3d_cuda_ctx *ctx, int exact);
SIFT3D_CUDA_EXPORT int sift3d_cuda_get_exact(sift3d_cuda_ctx *ctx);

/* Pinned host memory for image buffers (so the H2D copy in detect runs at PCIe
 * speed). Works without a context. */
SIFT3D_CUDA_EXPORT void *sift3d_cuda_host_alloc(size_t bytes);
SIFT3D_CUDA_EXPORT void sift3d_cuda_host_free(void *p);

/* (Re)allocate the device pyramid for this geometry. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_plan(sift3d_cuda_ctx *ctx, const sift3d_cuda_plan_desc *desc);

/* Run scale -> pyramid -> DoG max -> extrema -> ordering -> orientation.
 * image: nx*ny*nz float32, x fastest; on the host unless image_on_device != 0.
 * units: the units of THIS image (octave 0 uses them; octaves >= 1 use the plan's,
 * mirroring sift.c:651-656). Results stay on the device until fetched. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_detect(sift3d_cuda_ctx *ctx, const float *image,
                                          int image_on_device, const double units[3],
                                          double peak_thresh, double corner_thresh,
                                          int *num_candidates);

/* ---- one tall volume over several GPUs (z-slabs) -----------------------------------
 * The reference has no multi-device path; this is the decomposition of the same
 * build_gpyr / detect_extrema loops (sift.c:662-871) over z. Each rank plans a context
 * for its slab of planes plus halo planes on the cut faces, and owns a contiguous plane
 * range of every sharded octave. A blur only writes the owned planes of a level; the
 * halo planes are then filled from the neighbours through `exchange`, so every rank holds
 * exactly the values a single device would hold there. Octaves too small to shard are
 * run as a separate small context fed by sift3d_cuda_detect_from_seed. */
typedef struct {
    int num_sharded_octaves;            /* octaves 0 .. K-1 of the plan run on this slab */
    int own0[SIFT3D_CUDA_MAX_OCTAVES];  /* owned planes [own0[o], own1[o]) of octave o,  */
    int own1[SIFT3D_CUDA_MAX_OCTAVES];  /* in slab-array coordinates; entry K describes  */
                                        /* the seed level (K, -1) when the plan has it   */
    int org[SIFT3D_CUDA_MAX_OCTAVES];   /* plane of the whole volume's octave o that the  */
                                        /* slab array's plane 0 is: interpolation         */
                                        /* coordinates are formed in volume coordinates   */
                                        /* so that they round as on a single device       */
    int vol_nz[SIFT3D_CUDA_MAX_OCTAVES];/* planes of the whole volume's octave o (0: unknown; */
                                        /* a face of the array that is a face of the volume   */
                                        /* may then not lie inside a filter's reach of the    */
                                        /* owned range)                                       */
} sift3d_cuda_slab_desc;

typedef struct {
    void *user;
    /* element-wise max of `count` device floats over all ranks, in place */
    int (*reduce_max)(void *user, float *d_vals, int count);
    /* level (o, s) (s = -1: the seed) now holds its owned planes [own0, own1): fill the
     * other planes of the nx*ny*nz device array from the ranks that own them. Called
     * with the context's stream drained. */
    int (*exchange)(void *user, int o, int s, float *d_level, int nx, int ny, int nz,
                    int own0, int own1);
    /* != 0: the callbacks only ENQUEUE work on the context's stream (sift3d_cuda_ctx_stream),
     * e.g. peer copies behind event waits: detect does not drain the stream before calling
     * them (sift3d_cuda_multi, below). 0: they are called with the stream drained. */
    int in_stream;
    /* in_stream only: planes of the input image around the owned range that the first blur
     * reads; only those are uploaded from a host image (0: the whole slab array) */
    int input_reach;
} sift3d_cuda_slab_callbacks;

/* Returned by sift3d_cuda_detect_slab with in-stream callbacks when a candidate region was too
 * small: the regions have been resized, run the detect again on EVERY slab. */
#define SIFT3D_CUDA_SLAB_RETRY 2
/* As sift3d_cuda_detect for a z-slab. Candidates are those of the owned planes, with
 * slab-array coordinates; the caller offsets z and merges ranks in (o, s, rank) order. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_detect_slab(sift3d_cuda_ctx *ctx, const float *image,
                                               int image_on_device, const double units[3],
                                               double peak_thresh, double corner_thresh,
                                               const sift3d_cuda_slab_desc *slab,
                                               const sift3d_cuda_slab_callbacks *cb,
                                               int *num_candidates);

/* The context's stream as a plain pointer (a cudaStream_t), for in-stream slab callbacks. */
SIFT3D_CUDA_EXPORT void *sift3d_cuda_ctx_stream(sift3d_cuda_ctx *ctx);

/* As sift3d_cuda_detect, but `seed` already is level -1 of this context's octave 0 (the
 * 2x decimation sift.c:699-705 made in a larger pyramid): no im_scale, no first blur. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_detect_from_seed(sift3d_cuda_ctx *ctx, const float *seed,
                                                    int seed_on_device, const double units[3],
                                                    double peak_thresh, double corner_thresh,
                                                    int *num_candidates);

/* Octave number this context's octave 0 has in the pyramid it is a part of: reported in
 * keypoint.o by detect and subtracted from it by describe. Default 0. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_set_base_octave(sift3d_cuda_ctx *ctx, int base_octave);

/* Device address of level (o, s), s = -1 .. num_kp_levels + 1 (for the exchange
 * callbacks and for feeding a seed to another context). NULL if out of range. */
SIFT3D_CUDA_EXPORT float *sift3d_cuda_level_ptr(sift3d_cuda_ctx *ctx, int o, int s);

/* ---- one volume over several GPUs of ONE process (multi.cu) ---------------------------
 * The slab decomposition above with the exchange done by the library itself: one context and
 * one host thread per device, halo planes moved by peer copies (NVLink) that are enqueued
 * behind events -- no host round trip per level -- and only as many planes as the consumers of
 * a level read. The host library (libsift3D.so: sift3d_detect_describe_sharded) computes the
 * layout and the per-slab plans and merges the per-device results. */
#define SIFT3D_CUDA_MULTI_MAX_DEVICES 16
typedef struct sift3d_cuda_multi sift3d_cuda_multi;
typedef struct {
    int num_devices;
    int devices[SIFT3D_CUDA_MULTI_MAX_DEVICES];
    int nx, ny, nz;                 /* the whole volume */
    double units[3];
    int sharded_octaves;            /* K: octaves 0 .. K-1 run on the slabs */
    /* planes of octave 0: device r owns [g0, g1) (multiples of 2^K) and holds [a0, a1) */
    int g0[SIFT3D_CUDA_MULTI_MAX_DEVICES], g1[SIFT3D_CUDA_MULTI_MAX_DEVICES];
    int a0[SIFT3D_CUDA_MULTI_MAX_DEVICES], a1[SIFT3D_CUDA_MULTI_MAX_DEVICES];
    /* [o]: planes of octave o (each side) that the orientation / descriptor windows of the
     * keypoint levels reach (sift.c:936, 1453-1461) */
    int window_planes[SIFT3D_CUDA_MAX_OCTAVES];
} sift3d_cuda_multi_desc;
typedef struct {
    float total_ms, slabs_ms, tail_ms;      /* host wall clock of the last run: detect phase */
    float describe_ms;                      /* ... and describe phase (total_ms includes it) */
    float max_detect_ms, max_describe_ms;   /* slowest part */
    float exchanged_mbytes;                 /* halo planes moved between devices */
} sift3d_cuda_multi_timings;
/* slab_plans[r]: plan of device r's array (nx, ny, a1-a0; K or K+1 octaves); tail_plan: the
 * octaves >= K of the whole volume (NULL if there are none). */
SIFT3D_CUDA_EXPORT sift3d_cuda_multi *sift3d_cuda_multi_create(const sift3d_cuda_multi_desc *desc,
                                                               const sift3d_cuda_plan_desc *slab_plans,
                                                               const sift3d_cuda_plan_desc *tail_plan);
SIFT3D_CUDA_EXPORT void sift3d_cuda_multi_destroy(sift3d_cuda_multi *m);
/* Phase 1: detect on the whole volume `host_image` (nx*ny*nz floats, page-locked for speed). */
SIFT3D_CUDA_EXPORT int sift3d_cuda_multi_detect(sift3d_cuda_multi *m, const float *host_image,
                                                double peak_thresh, double corner_thresh);
/* part = device index, or num_devices for the unsharded octaves: its candidates in volume
 * coordinates, the part's scan order. Returns their number; valid until the next detect. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_multi_results(sift3d_cuda_multi *m, int part,
                                                 const sift3d_cuda_keypoint **cands);
/* Phase 2: every part describes its VALID candidates (in the part's order) on its own device,
 * all parts at once; the j-th valid candidate of part p goes to row rows[p][j] of `desc`
 * (page-locked host memory or device memory reachable from every device). */
SIFT3D_CUDA_EXPORT int sift3d_cuda_multi_describe(sift3d_cuda_multi *m, float *desc,
                                                  const int *const *rows);
SIFT3D_CUDA_EXPORT int sift3d_cuda_multi_get_timings(sift3d_cuda_multi *m, sift3d_cuda_multi_timings *t);

/* All candidates in the reference's scan order (o, s, z, y, x), each with its
 * orientation result (valid flag + R). out has room for max entries. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_fetch_candidates(sift3d_cuda_ctx *ctx,
                                                    sift3d_cuda_keypoint *out, int max);
/* Zero-copy view of the same records, octave by octave (octave_index = 0 .. num_octaves-1 of
 * this context, ascending = scan order): when the candidate regions live in page-locked
 * host memory (the orientation kernel writes them there directly) *records points at the
 * octave's records and the count is returned; valid until the next detect on this context.
 * Returns -1 when the records are device resident (use sift3d_cuda_fetch_candidates). */
SIFT3D_CUDA_EXPORT int sift3d_cuda_candidates_view(sift3d_cuda_ctx *ctx, int octave_index,
                                                   const sift3d_cuda_keypoint **records);
/* Companion of the view: 8 bytes per candidate (same order) with what the survivor compaction
 * needs from EVERY candidate -- its strength and whether it survived -- so that the host reads
 * the 88-byte records of the survivors only (a quarter of the candidates). */
typedef struct {
    float strength;
    int valid;
} sift3d_cuda_candidate_flag;
SIFT3D_CUDA_EXPORT int sift3d_cuda_candidates_flags(sift3d_cuda_ctx *ctx, int octave_index,
                                                    const sift3d_cuda_candidate_flag **flags);

/* Descriptors for num keypoints on the pyramid left by the last detect.
 * desc: num * 768 float32 on the host (or device if desc_on_device). */
SIFT3D_CUDA_EXPORT int sift3d_cuda_describe(sift3d_cuda_ctx *ctx,
                                            const sift3d_cuda_keypoint *keys, int num,
                                            float *desc, int desc_on_device);

/* As sift3d_cuda_describe, but the descriptor of keys[i] goes to row rows[i] of `desc` (which
 * must be device-accessible: device memory or page-locked host memory). Lets several
 * contexts fill disjoint rows of one result buffer (sift3d_cuda_multi). */
SIFT3D_CUDA_EXPORT int sift3d_cuda_describe_rows(sift3d_cuda_ctx *ctx,
                                                 const sift3d_cuda_keypoint *keys, int num,
                                                 float *desc, int desc_on_device, const int *rows);

/* Test / inspection hooks. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_level_info(sift3d_cuda_ctx *ctx, int o, int s, int dims[3],
                                              double units[3], double *scale);
SIFT3D_CUDA_EXPORT int sift3d_cuda_fetch_level(sift3d_cuda_ctx *ctx, int o, int s, float *host_out);
SIFT3D_CUDA_EXPORT int sift3d_cuda_dogmax(sift3d_cuda_ctx *ctx, int o, int s, float *out);
SIFT3D_CUDA_EXPORT int sift3d_cuda_blur(sift3d_cuda_ctx *ctx, const float *host_src,
                                        float *host_dst, int nx, int ny, int nz,
                                        const double units[3], const sift3d_cuda_filter *f,
                                        int force_generic);
/* One blur that allocates NOTHING. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it). Same kernels and same dispatch as
 * sift3d_cuda_blur (fused single-launch kernel where eligible, else the generic x / y / z
 * passes), but the two scratch volumes of the generic passes and the device copy of the taps
 * live in a file-scope `static __device__` array of the library: 2 x
 * SIFT3D_CUDA_STATIC_BLUR_MAX_VOXELS floats + 1 KiB = 8 MiB, held by every GPU that loads the
 * library until the process exits. Volumes with more voxels are rejected (-1, message on
 * stderr). The call issues no cudaMalloc / cudaFree / allocating library routine.
 *   d_src, d_dst  nx*ny*nz float32, x fastest, DEVICE-ACCESSIBLE and distinct (device memory, or
 *                 page-locked host memory of sift3d_cuda_host_alloc through its unified address)
 *   exact         arithmetic mode as sift3d_cuda_set_exact
 *   stream        a cudaStream_t passed as a plain pointer; NULL = the default stream
 *   sync          != 0: wait for completion before returning; 0: return once enqueued
 * Every call on a device shares the one scratch array, which keeps the last call's contents
 * (nothing in it has to start at zero: each call rewrites what it reads): calls that overlap
 * on the GPU -- different streams, threads or processes' worth of work on one device -- RACE.
 * Serialise them on one stream or with events. */
#define SIFT3D_CUDA_STATIC_BLUR_MAX_VOXELS (1 << 20)
SIFT3D_CUDA_EXPORT int sift3d_cuda_blur_static(int device, const float *d_src, float *d_dst,
                                               int nx, int ny, int nz, const double units[3],
                                               const sift3d_cuda_filter *f, int exact,
                                               void *stream, int sync);
/* Host-only: the work table of the fused unit-spacing blur (no device needed). `tiles` column
 * tiles of `nzr` output planes each are cut into `ncta` contiguous ranges of the flattened
 * tile-major space; bounds[0..ncta] receives the range limits. Returns the cost (output planes
 * + warm-up planes) of the most expensive range, or -1 on invalid arguments. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_blur_chunk_table(int ncta, int tiles, int nzr, int hw,
                                                    int top_is_volume_top, int *bounds);
/* Host-only: the banded weights of one pass of the fast-mode blur for a general tap spacing (no
 * device needed; blur_band.cuh). A line of `n_line` samples of which the array holds
 * [org, n_line) (org > 0: the lower cut face of a z-slab), tap spacing `unit_factor` (1 / units
 * of the axis, as a float, like imutil.c:760), filter `f`. Position p of the array gets
 * out[p] = sum_j w[j * n + p] * in[base[p] + j], j < *band, n = n_line - org: the same samples the
 * reference's coordinate rule reads (imutil.c:742-861), products of one sample folded together.
 * `w` holds w_capacity floats, `base` n ints. Returns 0, 1 if the band does not fit w_capacity or
 * the line is not representable, -1 on invalid arguments. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_blur_band_table(int n_line, int org, float unit_factor,
                                                   const sift3d_cuda_filter *f, float *w,
                                                   size_t w_capacity, int *base, int *band);
SIFT3D_CUDA_EXPORT int sift3d_cuda_get_timings(sift3d_cuda_ctx *ctx, sift3d_cuda_timings *t);
/* Number of kernel launches issued by this context since creation. */
SIFT3D_CUDA_EXPORT uint64_t sift3d_cuda_launch_count(sift3d_cuda_ctx *ctx);
SIFT3D_CUDA_EXPORT int sift3d_cuda_synchronize(sift3d_cuda_ctx *ctx);

/* Device-resident copies of image buffers (so a caller can keep inputs in HBM and
 * skip the per-call upload). Returns NULL / -1 on failure. */
SIFT3D_CUDA_EXPORT void *sift3d_cuda_device_upload(int device, const float *host, size_t count);
SIFT3D_CUDA_EXPORT void sift3d_cuda_device_free(int device, void *dptr);

/* Asynchronous upload on a copy stream of its own, so that the transfer of the NEXT volume
 * overlaps the kernels of the current one (the reference reads its input from host memory,
 * sift.c:629-659; on a GPU the PCIe copy is a fifth of the step and hides completely).
 * begin() returns at once; `host` must be page-locked (sift3d_cuda_host_alloc) and unchanged
 * until the transfer is consumed. `reuse` (may be NULL) recycles a previous ticket's device
 * buffer. wait() makes the context's stream wait for the transfer without blocking the host. */
typedef struct sift3d_cuda_upload sift3d_cuda_upload;
SIFT3D_CUDA_EXPORT sift3d_cuda_upload *sift3d_cuda_upload_begin(int device, const float *host,
                                                                size_t count,
                                                                sift3d_cuda_upload *reuse);
SIFT3D_CUDA_EXPORT const float *sift3d_cuda_upload_ptr(const sift3d_cuda_upload *up);
SIFT3D_CUDA_EXPORT int sift3d_cuda_upload_device(const sift3d_cuda_upload *up);
SIFT3D_CUDA_EXPORT int sift3d_cuda_upload_wait(sift3d_cuda_ctx *ctx, sift3d_cuda_upload *up);
/* Block the host until the transfer has completed. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_upload_sync(sift3d_cuda_upload *up);
SIFT3D_CUDA_EXPORT void sift3d_cuda_upload_free(sift3d_cuda_upload *up);
SIFT3D_CUDA_EXPORT int sift3d_cuda_ctx_device(sift3d_cuda_ctx *ctx);

/* User timing marks on the context's stream (slots 0..7): record now / elapsed ms
 * between two recorded marks (synchronises on `b`). */
SIFT3D_CUDA_EXPORT int sift3d_cuda_mark(sift3d_cuda_ctx *ctx, int slot);
SIFT3D_CUDA_EXPORT int sift3d_cuda_elapsed_ms(sift3d_cuda_ctx *ctx, int a, int b, float *ms);

/* Per-blur profiling: when enabled, detect brackets every blur (all launches that
 * produce one Gaussian level) with CUDA events. blur_ms returns the duration of the blur
 * that produced level (o, s) in the last detect, and the number of launches it took. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_set_profiling(sift3d_cuda_ctx *ctx, int enabled);
/* Start and end of the blur of level (o, s) relative to the start of the last detect, in
 * milliseconds (profiling mode): the timeline of the octave streams. */
SIFT3D_CUDA_EXPORT int sift3d_cuda_level_span_ms(sift3d_cuda_ctx *ctx, int o, int s, float *begin_ms,
                                                 float *end_ms);
SIFT3D_CUDA_EXPORT int sift3d_cuda_blur_ms(sift3d_cuda_ctx *ctx, int o, int s, float *ms,
                                           int *launches);

#ifdef __cplusplus
}
#endif

#endif /* SIFT3D_CUDA_H */
