/* sift3d.h -- public C API of the B200-native SIFT3D detect + describe path.
 *
 * This is the drop-in boundary: the 27 functions exported by the reference's
 * libsift3D.so.2 (fatimp/SIFT3D v2.0), with the same names, argument order,
 * opaque handle types, return codes and ownership rules, so that a program
 * written against the reference's <sift.h>/<imutil.h>/<imtypes.h> relinks
 * against this library unchanged. Citations are to /root/reference/sift3d/.
 *
 *   types, return codes        imtypes.h:14-44
 *   detector + setters         sift.h:24-78      (impl sift.c:499-565, 1841-1859)
 *   detect / extract           sift.h:92-111     (impl sift.c:1217-1249, 1615-1635)
 *   keypoint store             sift.h:121-165    (impl sift.c:1644-1671, 1741-1803, 1861-1900)
 *   descriptor store           sift.h:175-208    (impl sift.c:1683-1726, 1807-1830, 1873-1883)
 *   images                     imutil.h:39-65    (impl imutil.c:1639-1674)
 *   matrices                   imutil.h:77-110   (impl imutil.c:1676-1710)
 *
 * All five handle types are opaque in the reference's public headers too, so the
 * layouts behind them are this library's own. The compute behind
 * sift3d_detect_keypoints / sift3d_extract_descriptors runs on the GPU through
 * libsift3d_cuda.so (include/sift3d_cuda.h); there is no CPU fallback.
 *
 * Additive extensions (not in the reference) are grouped at the end.
 */
#ifndef SIFT3D_B200_H
#define SIFT3D_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define SIFT3D_EXPORT __attribute__((visibility("default")))

/* Return codes (imtypes.h:21-25) */
#define SIFT3D_SUCCESS 0
#define SIFT3D_FAILURE -1
#define SIFT3D_TRUE 1
#define SIFT3D_FALSE 0
/* sift3d_read_image() status values (imutil.h:17-23) */
#define SIFT3D_UNSUPPORTED_FILE_TYPE 2
#define SIFT3D_WRAPPER_NOT_COMPILED 3

typedef struct _sift3d_detector sift3d_detector;
typedef struct _sift3d_keypoint_store sift3d_keypoint_store;
typedef struct _sift3d_descriptor_store sift3d_descriptor_store;
typedef struct _sift3d_image sift3d_image;
typedef struct _sift3d_mat_rm sift3d_mat_rm;

/* Element type of a matrix (imtypes.h:40-44) */
typedef enum { SIFT3D_DOUBLE, SIFT3D_FLOAT, SIFT3D_INT } sift3d_mat_type;

/* ---- detector ------------------------------------------------------------------- */

/* Detector with default parameters (peak 0.1, corner 0.4, 3 keypoint levels,
 * sigma_n 1.15, sigma0 1.6; sift.c:31-35). NULL on failure, including "no CUDA
 * device". Free with sift3d_free_detector(). */
SIFT3D_EXPORT sift3d_detector *sift3d_make_detector(void);
SIFT3D_EXPORT void sift3d_free_detector(sift3d_detector *);

/* peak_thresh in (0, 1] (sift.c:499-509) */
SIFT3D_EXPORT int sift3d_detector_set_peak_thresh(sift3d_detector *const, const double);
/* corner_thresh in [0, 1] (sift.c:512-523) */
SIFT3D_EXPORT int sift3d_detector_set_corner_thresh(sift3d_detector *const, const double);
/* keypoint levels per octave (sift.c:527-533) */
SIFT3D_EXPORT int sift3d_detector_set_num_kp_levels(sift3d_detector *const, const unsigned int);
/* nominal scale of the input, >= 0 (sift.c:537-549) */
SIFT3D_EXPORT int sift3d_detector_set_sigma_n(sift3d_detector *const, const double);
/* scale of level 0 of octave 0, >= 0 (sift.c:553-565) */
SIFT3D_EXPORT int sift3d_detector_set_sigma0(sift3d_detector *const, const double);

/* Keypoint positions + orientations of a single-channel image. The image is
 * consumed during the call and may be freed right after it (sift.h:83-95). */
SIFT3D_EXPORT int sift3d_detect_keypoints(sift3d_detector *const detector,
                                          const sift3d_image *const image,
                                          sift3d_keypoint_store *const store);

/* Descriptors at the keypoints of `kp_store`, on the pyramid left in `detector`
 * by its last sift3d_detect_keypoints() (sift.h:97-111). */
SIFT3D_EXPORT int sift3d_extract_descriptors(sift3d_detector *const detector,
                                             const sift3d_keypoint_store *const kp_store,
                                             sift3d_descriptor_store *const desc_store);

/* ---- keypoint store -------------------------------------------------------------- */

SIFT3D_EXPORT sift3d_keypoint_store *sift3d_make_keypoint_store(void);
SIFT3D_EXPORT void sift3d_free_keypoint_store(sift3d_keypoint_store *);
/* N x 3 double, base-octave coordinates (sift.c:1644-1671) */
SIFT3D_EXPORT int sift3d_keypoint_store_to_mat_rm(const sift3d_keypoint_store *const,
                                                  sift3d_mat_rm *const);
/* CSV (".gz" => gzip): strength,x,y,z,o,sd,R00..R22 (sift.c:1741-1803) */
SIFT3D_EXPORT int sift3d_keypoint_store_save(const char *path,
                                             const sift3d_keypoint_store *const);
/* Descending strength; keep at most `limit` if limit != 0 (sift.c:1885-1900) */
SIFT3D_EXPORT void sift3d_keypoint_store_sort_by_strength(sift3d_keypoint_store *const,
                                                          int limit);

/* ---- descriptor store ------------------------------------------------------------ */

SIFT3D_EXPORT sift3d_descriptor_store *sift3d_make_descriptor_store(void);
SIFT3D_EXPORT void sift3d_free_descriptor_store(sift3d_descriptor_store *);
SIFT3D_EXPORT int sift3d_descriptor_store_save(const char *path,
                                               const sift3d_descriptor_store *const);
/* N x 771 float: x, y, z, then 768 elements ordered hist*12 + bin (sift.c:1683-1726) */
SIFT3D_EXPORT int sift3d_descriptor_store_to_mat_rm(const sift3d_descriptor_store *const,
                                                    sift3d_mat_rm *const);

/* ---- images ------------------------------------------------------------------------ */

/* nx*ny*nz*nc float32, zero-filled, x fastest (imutil.h:58-63). The buffer is
 * page-locked so sift3d_detect_keypoints() uploads it at PCIe speed. */
SIFT3D_EXPORT sift3d_image *sift3d_make_image(const int nx, const int ny, const int nz,
                                              const int nc);
SIFT3D_EXPORT void sift3d_free_image(sift3d_image *);
/* NIfTI-1 (.nii / .nii.gz); NULL on failure (imutil.h:48-55) */
SIFT3D_EXPORT sift3d_image *sift3d_read_image(const char *path);
SIFT3D_EXPORT float *sift3d_image_data(const sift3d_image *);

/* ---- matrices ---------------------------------------------------------------------- */

SIFT3D_EXPORT sift3d_mat_rm *sift3d_make_mat_rm(void);
SIFT3D_EXPORT void sift3d_free_mat_rm(sift3d_mat_rm *);
SIFT3D_EXPORT void *sift3d_mat_rm_data(sift3d_mat_rm *);
SIFT3D_EXPORT void sift3d_mat_rm_dimensions(const sift3d_mat_rm *, int *num_cols, int *num_rows);
SIFT3D_EXPORT sift3d_mat_type sift3d_mat_rm_type(const sift3d_mat_rm *);

/* ---- extensions (additive; nothing above depends on them) ------------------------- */

/* Voxel spacing. The reference can only set units through its NIfTI reader
 * (nifti.c:89-91); programs that build images in memory need this. */
SIFT3D_EXPORT int sift3d_image_set_units(sift3d_image *, double ux, double uy, double uz);
SIFT3D_EXPORT void sift3d_image_get_units(const sift3d_image *, double *ux, double *uy, double *uz);
SIFT3D_EXPORT void sift3d_image_dimensions(const sift3d_image *, int *nx, int *ny, int *nz, int *nc);

/* Keep a copy of the image in the memory of CUDA device `device` (snapshot of the
 * current host contents); detectors on that device then skip the upload. device < 0
 * drops the copy. */
SIFT3D_EXPORT int sift3d_image_make_resident(sift3d_image *, int device);

/* Start an asynchronous copy of the current contents to CUDA device `device` and return at
 * once; the next sift3d_detect_keypoints on a detector of that device consumes it without
 * blocking the host on the transfer. Lets the upload of the next volume overlap the kernels
 * of the current one. The host buffer must stay unchanged until that detect. */
SIFT3D_EXPORT int sift3d_image_prefetch(sift3d_image *, int device);
/* Block until that copy has completed (not needed before detect). */
SIFT3D_EXPORT int sift3d_image_prefetch_wait(sift3d_image *);

/* Detector bound to a given CUDA device (sift3d_make_detector() uses
 * $SIFT3D_CUDA_DEVICE or device 0). */
SIFT3D_EXPORT sift3d_detector *sift3d_make_detector_on_device(int device);

/* Extension: fill a sift3d_cuda_plan_desc (sift3d_cuda.h) with the scales and filters this
 * detector's parameters give for the octaves base_octave .. base_octave + num_octaves - 1
 * of a pyramid, whose first described octave is nx x ny x nz with the given units
 * (num_octaves 0: as many as fit, sift.c:441-452). Used by the z-slab multi-GPU harness,
 * which drives the CUDA contexts itself. */
SIFT3D_EXPORT int sift3d_detector_fill_plan(const sift3d_detector *, int nx, int ny, int nz,
                                            double ux, double uy, double uz, int base_octave,
                                            int num_octaves, void *plan_desc);

/* im_write (imutil.c:385-402, not exported by the fork): float32 NIfTI-1, .nii or .nii.gz. */
SIFT3D_EXPORT int sift3d_write_image(const char *path, const sift3d_image *);

/* The NIfTI-1 / Analyze reader and writer behind sift3d_read_image / sift3d_write_image
 * without the page-locked image object (usable without a GPU). dims = nx, ny, nz, nc; data is
 * channel-fastest, c + nc*(x + nx*(y + ny*z)); *data is released with sift3d_nifti_free. */
SIFT3D_EXPORT int sift3d_nifti_read_raw(const char *path, int dims[4], double units[3], float **data);
SIFT3D_EXPORT int sift3d_nifti_write_raw(const char *path, const int dims[4], const double units[3],
                                         const float *data);
SIFT3D_EXPORT void sift3d_nifti_free(float *data);

/* Number of keypoints / descriptors held by a store. */
SIFT3D_EXPORT int sift3d_keypoint_store_count(const sift3d_keypoint_store *);
SIFT3D_EXPORT int sift3d_descriptor_store_count(const sift3d_descriptor_store *);

/* One keypoint record: o, s, octave-local x/y/z, scale, strength and the row-major
 * rotation matrix. Returns SIFT3D_FAILURE if index is out of range. */
SIFT3D_EXPORT int sift3d_keypoint_store_get(const sift3d_keypoint_store *, int index, int *o,
                                            int *s, double xyz[3], double *sd, float *strength,
                                            float R[9]);

/* Batch: detect + describe `num` volumes over `num_det` detectors, each
 * bound to a device of its own (sift3d_make_detector_on_device). Volume i is processed by
 * detector i % num_det exactly as sift3d_detect_keypoints + sift3d_extract_descriptors
 * (sift.h:92-111) would; every detector is driven by a host thread of its own and the upload of
 * its next volume overlaps the kernels of the current one. descs may be NULL (detection only)
 * and so may single entries of it; a volume without keypoints leaves its descriptor store
 * empty. Image buffers must stay unchanged during the call. Returns SIFT3D_FAILURE if any
 * volume failed. */
SIFT3D_EXPORT int sift3d_detect_describe_batch(sift3d_detector *const *dets, int num_det,
                                               sift3d_image *const *ims,
                                               sift3d_keypoint_store *const *kps,
                                               sift3d_descriptor_store *const *descs, int num);

/* One tall volume over several devices: z-slabs of the volume, one
 * per device, with halo exchange over NVLink between the slabs -- the same result as
 * sift3d_detect_keypoints + sift3d_extract_descriptors on one device (the octave loops of
 * sift.c:662-871, 1109-1167, 1442-1536 decomposed over z). `sharded_octaves` octaves run on
 * the slabs, the smaller ones on devices[0]. desc may be NULL (detection only). The detector
 * only lends its parameters; its own pyramid is not touched. */
SIFT3D_EXPORT int sift3d_detect_describe_sharded(sift3d_detector *, const sift3d_image *,
                                                 const int *devices, int num_devices,
                                                 int sharded_octaves, sift3d_keypoint_store *,
                                                 sift3d_descriptor_store *);
/* total, slab detects, unsharded octaves' detect, describe phase, slowest slab detect, slowest
 * part's describe (ms of host wall clock); MiB of halo planes moved between devices */
SIFT3D_EXPORT int sift3d_sharded_timings(sift3d_detector *, float out[7]);
/* The layout rule (host only): owned [g0, g1) / held [a0, a1) planes per device; returns the
 * halo in planes of octave 0, or -1 if nz cannot be split into num slabs aligned to 2^K. */
SIFT3D_EXPORT int sift3d_sharded_layout(int nz, int num, int K, int window_planes, int max_half_width,
                                        double uz, int *g0, int *g1, int *a0, int *a1);

/* Upstream-cased aliases (bbrister/SIFT3D spells the path SIFT3D_detect_keypoints /
 * SIFT3D_extract_descriptors on Image / Keypoint_store / SIFT3D_Descriptor_store; the fork
 * mounted as the reference renamed them, sift.h:92-111). Same argument order and return
 * codes; the types are the fork's opaque handles under the upstream names. */
typedef sift3d_detector SIFT3D;
typedef sift3d_image Image;
typedef sift3d_keypoint_store Keypoint_store;
typedef sift3d_descriptor_store SIFT3D_Descriptor_store;
SIFT3D_EXPORT int SIFT3D_detect_keypoints(SIFT3D *, const Image *, Keypoint_store *);
SIFT3D_EXPORT int SIFT3D_extract_descriptors(SIFT3D *, const Keypoint_store *, SIFT3D_Descriptor_store *);

/* The inner CUDA context of a detector (include/sift3d_cuda.h), for timing and
 * pyramid inspection by tests and benchmarks. */
SIFT3D_EXPORT void *sift3d_detector_cuda_ctx(sift3d_detector *);

#ifdef __cplusplus
}
#endif

#endif /* SIFT3D_B200_H */
