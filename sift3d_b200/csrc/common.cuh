// common.cuh -- shared declarations of the CUDA layer (libsift3d_cuda.so).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "../../include/sift3d_cuda.h"

#define S3D_ERR(...) fprintf(stderr, __VA_ARGS__)

#define S3D_CK(call)                                                               \
    do {                                                                           \
        cudaError_t e_ = (call);                                                   \
        if (e_ != cudaSuccess) {                                                   \
            S3D_ERR("sift3d_cuda: %s failed: %s (%s:%d)\n", #call,                  \
                    cudaGetErrorString(e_), __FILE__, __LINE__);                   \
            return -1;                                                             \
        }                                                                          \
    } while (0)

namespace s3d {

constexpr int kNumSMs = 148;  // B200

// A pyramid level resident in HBM: dense float32, x fastest (immacros.h:58-59).
struct Level {
    float *d = nullptr;
    int nx = 0, ny = 0, nz = 0;
    double ux = 1, uy = 1, uz = 1;
    double scale = 0;
    size_t nvox() const { return (size_t) nx * ny * nz; }
};

// Taps passed by value in the kernel parameter bank (constant memory).
// Function attributes and __constant__ symbols belong to a device: one-time set-up is done once
// PER DEVICE (a process may hold detectors on several GPUs, sift3d_detector_set_device).
constexpr int kMaxDevices = 64;
struct PerDeviceOnce {
    std::mutex mu;
    bool done[kMaxDevices] = {};
    // Runs f() (0 = success) the first time it is called on the current device; a second host
    // thread driving the same device waits until the set-up has been made (it would otherwise
    // launch before the kernel's shared-memory attribute is in place).
    template <typename F>
    int run(F f) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return f();
        std::lock_guard<std::mutex> lock(mu);
        if (done[dev]) return 0;
        const int rc = f();
        if (rc == 0) done[dev] = true;
        return rc;
    }
};

constexpr int kMaxParamTaps = 33;  // half-width <= 16
struct TapsParam {
    float k[kMaxParamTaps];
};

// Per-axis sampling parameters of convolve_sep_gen (imutil.c:748-763).
struct AxisSpec {
    int n;        // samples along the axis
    float uf;     // unit_factor = (float)(unit / units[dim])
    int uhw;      // (int) ceilf(half_width * unit_factor)
    int sh;       // unit_factor == 2^-sh exactly (dyadic fast path), else -1
};

// Candidate extremum as written by the extrema kernel.
struct Cand {
    unsigned long long key;  // s << 60 | z << 40 | y << 20 | x
    float strength;
    int pad;
};

// ---- kernels / launchers implemented in the .cu files -----------------------------------

// blur.cu
// SMs the fused one-wave blur leaves free on this thread's following launches: while other
// octaves work on their own (higher-priority) streams a 148-CTA persistent kernel could not
// place all its CTAs at once and the stragglers would double its time.
void set_fused_sm_reserve(int sms);
int launch_blur(cudaStream_t st, const float *src, float *dst, float *tmp_a, float *tmp_b,
                int nx, int ny, int nz, const double units[3], const sift3d_cuda_filter &f,
                const float *d_taps, const float *d_divisor, float *d_dogmax_slot,
                const float *dog_prev, float *ds_dst, bool force_generic, uint64_t *launches,
                int z0 = 0, int z1 = -1,   // output planes [z0, z1); z1 < 0: the whole volume
                int zorg = 0,              // z-slab: volume plane of array plane 0
                bool exact = true,         // 0-ULP arithmetic (else one FFMA per tap, see blur_zmarch.cuh)
                int vol_nz = 0,            // z-slab: planes of the whole volume at this octave (0: unknown)
                bool no_alloc = false);    // never allocate (no band tables: literal kernels for general spacings)
// Chain of tiny octaves in one launch (blur.cu: tiny_octaves_kernel)
constexpr int kTinyThreads = 1024;
constexpr int kTinyMaxVox = 4096;   // voxels of the largest octave the chain kernel takes
constexpr int kTinyMaxOct = 8;
struct TinyOctave {
    int nx, ny, nz;
    float uf[3];                                  // unit factor per axis (1 / level units)
    int sh[3];                                    // dyadic exponent per axis or -1
    float *lvl[SIFT3D_CUDA_MAX_FILTERS + 1];      // level -1 .. nfilt-1
    float *dogmax[SIFT3D_CUDA_MAX_FILTERS];       // [s]: slot for max|G(s-1) - G(s)| or nullptr
};
// Fast mode: one pass of a blur as banded multiply-adds (blur_band.cuh, positions in blocks of 4):
// out[p] = sum_r w[(r * nblk + p / 4) * 4 + p % 4] * in[bb[p / 4] + r], r < BB.
struct TinyBand {
    const float *w;
    const int *bb;
    int BB, nblk;
};
struct TinyChainParams {
    int noct, nfilt, ds_level;
    int banded;                                   // fast mode: `band` is filled, the literal taps are not used
    TinyBand band[kTinyMaxOct][SIFT3D_CUDA_MAX_FILTERS][3];
    const float *taps[SIFT3D_CUDA_MAX_FILTERS];   // filter s: level s-1 -> s
    int hw[SIFT3D_CUDA_MAX_FILTERS];
    TinyOctave oct[kTinyMaxOct];
};
int launch_tiny_octaves(cudaStream_t st, const TinyChainParams &P, uint64_t *launches);
void tiny_axis(int n, double unit, float *uf, int *sh);
// Host: band of one axis of a tiny octave (0: filled, 1: not representable, -1: error)
int tiny_band(int n, double unit, const sift3d_cuda_filter &f, TinyBand *out);
int launch_downsample(cudaStream_t st, const float *src, float *dst, int snx, int sny, int snz,
                      uint64_t *launches);
int launch_absmax(cudaStream_t st, const float *src, size_t n, float *d_out, uint64_t *launches);
int launch_absdiffmax(cudaStream_t st, const float *a, const float *b, size_t n, float *d_out,
                      uint64_t *launches);

// extrema.cu
int launch_extrema(cudaStream_t st, const float *const *g, int num_kp_levels, int nx, int ny,
                   int nz, const float *d_dogmax, double peak_thresh, Cand *d_cand, int cap,
                   int *d_count, uint64_t *launches, int z0 = 0, int z1 = -1,
                   int s_begin = 0, int s_end = -1)  /* keypoint levels [s_begin, s_end) */;  // planes [z0, z1)
// all octaves in one launch; counts stay on the device
struct OrderScratch {   // for the bucket path of busy regions (extrema.cu)
    Cand *grouped;      // as large as the candidate regions
    int *ints;          // 3 * nbkt[o] ints per octave at off[o]
    int nz[SIFT3D_CUDA_MAX_OCTAVES], nbkt[SIFT3D_CUDA_MAX_OCTAVES];
    size_t off[SIFT3D_CUDA_MAX_OCTAVES];
};
int launch_order_all(cudaStream_t st, const Cand *d_raw, Cand *d_sorted, const int *base,
                     const int *cap, int noct, const int *d_counts, uint64_t *launches,
                     int o_begin = 0, int o_end = -1,   // octaves [o_begin, o_end) only
                     const OrderScratch *scratch = nullptr);

// orient.cu
struct OctaveInfo {  // what the per-candidate kernels need to know about one octave
    const float *lvl[SIFT3D_CUDA_MAX_FILTERS];  // Gaussian level of keypoint level s (gpyr s)
    double sd[SIFT3D_CUDA_MAX_FILTERS];         // level scale (imutil.c:1578-1579)
    int o;
    int nx, ny, nz;
    float ux, uy, uz;  // level units narrowed to float (sift.c:88-90)
};
// One voxel of an orientation window: integer offset from the candidate and its weight. The
// window of a candidate is a sphere around INTEGER coordinates (sift.c:1124), so offsets and
// weights are the same for every candidate of a keypoint level and are tabulated once.
struct OrientEntry {
    int dx, dy, dz;
    float w;
};
// Everything the orientation kernel needs to walk the candidate regions of all octaves.
struct OrientPlan {
    const OrientEntry *tab[SIFT3D_CUDA_MAX_OCTAVES][SIFT3D_CUDA_MAX_FILTERS];  // window tables, or all NULL
    const int *ntab;  // entries of tab[o][s] at [o * SIFT3D_CUDA_MAX_FILTERS + s]
    OctaveInfo oct[SIFT3D_CUDA_MAX_OCTAVES];
    int base[SIFT3D_CUDA_MAX_OCTAVES];  // first slot of the octave's candidate region
    int cap[SIFT3D_CUDA_MAX_OCTAVES];
    int noct;
};
// One persistent launch over the candidates of every octave (counts read on the device,
// work handed out through *d_work, which must be zero). Output slot = rank of the
// candidate in the concatenated (o, s, z, y, x) order.
// Fill the window tables of plan `h` (host copy: octave geometry and level scales filled in):
// `d_entries` has room for orient_table_capacity(h, kp) entries; sets h.tab / h.ntab.
size_t orient_table_capacity(const OrientPlan &h, int kp, size_t off[SIFT3D_CUDA_MAX_OCTAVES][SIFT3D_CUDA_MAX_FILTERS]);
int launch_orient_tables(cudaStream_t st, OrientPlan &h, int kp, OrientEntry *d_entries, int *d_counts,
                         uint64_t *launches);
int launch_orient_all(cudaStream_t st, const Cand *d_sorted, const OrientPlan *d_plan,
                      const int *d_counts, int *d_work, double corner_thresh,
                      sift3d_cuda_keypoint *d_out, uint64_t *launches,
                      int o_begin = 0, int o_end = -1,   // octaves [o_begin, o_end); own d_work slot
                      sift3d_cuda_candidate_flag *d_flags = nullptr,  // (strength, valid) per slot, or NULL
                      bool shared_device = false);  // other detectors run on this device: small grids

// descriptor.cu
struct DescJob {
    const float *img;
    int nx, ny, nz;
    float ux, uy, uz;
    float cx, cy, cz;
    float sigma;         // (float)(sd * desc_sig_fctr)  (sift.c:1453)
    float R[9];
    int out_index;       // descriptor slot (jobs are launched largest window first)
    int max_rows;        // upper bound of (y, z) rows of the window (host estimate)
};
int upload_mesh(void);
// jobs may be reordered by the caller; max_rows = max over jobs of DescJob::max_rows
int launch_descriptors(cudaStream_t st, const DescJob *d_jobs, int n, int max_rows, float *d_desc,
                       uint64_t *launches);
// copy n 32-bit words from page-locked host memory to device memory with a kernel
int launch_pull_words(cudaStream_t st, const void *pinned_src, void *dst, size_t n, uint64_t *launches);
int desc_job_rows(const DescJob &J);  // host: upper bound of window rows of one job

}  // namespace s3d
