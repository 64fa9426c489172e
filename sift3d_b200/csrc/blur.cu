// blur.cu -- Gaussian scale-space blur, 2x decimation and the max-abs reductions.
//
// Reference semantics (paths under /root/reference/sift3d):
//   apply_Sep_FIR_filter  imutil.c:1127-1206   x pass, then y, then z, fp32 between passes
//   convolve_sep_gen      imutil.c:742-861     taps spaced unit/units[dim] voxels apart,
//                                              sampled by linear interpolation; interior
//                                              [uhw, n-2-uhw] without bounds logic, elsewhere
//                                              mirror-left / mirror-minus-0.1-right
//   im_downsample_2x      imutil.c:591-617     dst[x,y,z] = src[2x,2y,2z]
//   im_max_abs / im_scale imutil.c:681-713     max |v|, then v /= max (true division)
//   im_subtract           imutil.c:719-739     DoG = G(s) - G(s+1) (never materialised here)
//
// Bit parity: the reference build (x86-64, no -mfma) rounds every multiply and every add
// to fp32 separately and accumulates taps in ascending d. All arithmetic below uses
// __fmul_rn/__fadd_rn/__fsub_rn/__fdiv_rn so nvcc can never contract it into FMAs.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace s3d {

// ---------------------------------------------------------------------------------------
// Literal per-tap sampling (SAMP_AND_ACC, imutil.c:779-796).
// ---------------------------------------------------------------------------------------
// Coordinates are those of the whole line: `line` holds samples org .. n-1 of it (org > 0 only
// for the z axis of a z-slab, whose array starts at plane org of the volume -- the fractions
// then round exactly as on a single device).
// The sample type T is float (one line) or float4 (four x-adjacent lines of a y or z pass:
// coordinates, fractions and addresses are those of the line, so they are formed once for
// four voxels; every component sees exactly the scalar operation sequence).
template <typename T> struct Samp;
template <> struct Samp<float> {
    static __device__ __forceinline__ float zero() { return 0.0f; }
    template <bool ROWS>
    static __device__ __forceinline__ float ld(const float *p, size_t) { return *p; }
    static __device__ __forceinline__ float div(float a, float dv) { return __fdiv_rn(a, dv); }
    static __device__ __forceinline__ float lerp(float fr, float a, float b) {
        return __fadd_rn(__fmul_rn(__fsub_rn(1.0f, fr), a), __fmul_rn(fr, b));
    }
    static __device__ __forceinline__ float acc(float s, float tap, float v) {
        return __fadd_rn(s, __fmul_rn(tap, v));
    }
};
template <> struct Samp<float4> {
    static __device__ __forceinline__ float4 zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
    // ROWS: the four voxels are the same column of four lines `ls` apart (x pass over four
    // rows); else four adjacent columns (16-byte load)
    template <bool ROWS>
    static __device__ __forceinline__ float4 ld(const float *p, size_t ls) {
        if (ROWS) return make_float4(p[0], p[ls], p[2 * ls], p[3 * ls]);
        return *reinterpret_cast<const float4 *>(p);
    }
    static __device__ __forceinline__ float4 div(float4 a, float dv) {
        return make_float4(__fdiv_rn(a.x, dv), __fdiv_rn(a.y, dv), __fdiv_rn(a.z, dv), __fdiv_rn(a.w, dv));
    }
    static __device__ __forceinline__ float4 lerp(float fr, float4 a, float4 b) {
        const float om = __fsub_rn(1.0f, fr);
        return make_float4(__fadd_rn(__fmul_rn(om, a.x), __fmul_rn(fr, b.x)),
                           __fadd_rn(__fmul_rn(om, a.y), __fmul_rn(fr, b.y)),
                           __fadd_rn(__fmul_rn(om, a.z), __fmul_rn(fr, b.z)),
                           __fadd_rn(__fmul_rn(om, a.w), __fmul_rn(fr, b.w)));
    }
    static __device__ __forceinline__ float4 acc(float4 s, float tap, float4 v) {
        return make_float4(__fadd_rn(s.x, __fmul_rn(tap, v.x)), __fadd_rn(s.y, __fmul_rn(tap, v.y)),
                           __fadd_rn(s.z, __fmul_rn(tap, v.z)), __fadd_rn(s.w, __fmul_rn(tap, v.w)));
    }
};

template <typename T = float, bool ROWS = false>
__device__ __forceinline__ T samp(const float *__restrict__ line, size_t stride, float c,
                                  int n, bool scale, float dv, int org, size_t ls = 0) {
    int lo = (int) c;  // truncation toward zero, as the C cast
    const float fr = __fsub_rn(c, (float) lo);
    // Only reachable with lo outside [0, n-2] where the reference itself reads out of
    // bounds (single mirror on lines shorter than the kernel); keep the access legal.
    lo = max(org, min(lo, n - 2)) - org;
    T a = Samp<T>::template ld<ROWS>(line + (size_t) lo * stride, ls);
    T b = Samp<T>::template ld<ROWS>(line + (size_t) (lo + 1) * stride, ls);
    if (scale) {
        a = Samp<T>::div(a, dv);
        b = Samp<T>::div(b, dv);
    }
    return Samp<T>::lerp(fr, a, b);
}

// One output sample of convolve_sep_gen along a line (imutil.c:798-856).
template <typename T = float, bool ROWS = false>
__device__ __forceinline__ T conv_point_generic(const float *__restrict__ line,
                                                size_t stride, int p, int n,
                                                const float *__restrict__ taps, int hw,
                                                float uf, int uhw, bool scale, float dv,
                                                int sh = -1, int org = 0, size_t ls = 0) {
    using S = Samp<T>;
    const int start = uhw;
    const int end = n - 2 - uhw;
    const int dim_end = n - 1;
    T acc = S::zero();
    if (sh >= 0 && p >= start && p <= end) {
        // Interior pass with a dyadic unit factor 2^-sh: p - d*uf and the way back are exact,
        // so tap d samples position p + (-d >> sh) with fraction ((-d) mod 2^sh) * 2^-sh --
        // the same values the carried-coordinate loop below produces, without the
        // float <-> int round trips. frac == 0 taps are the sample itself: 1*a + 0*b == a.
        const float *__restrict__ q = line + (size_t) (p - org) * stride;
        const ptrdiff_t st = (ptrdiff_t) stride;
        for (int d = -hw; d <= hw; d++) {
            const float tap = taps[d + hw];
            const int off = (-d) >> sh;
            const int fi = (-d) & ((1 << sh) - 1);
            T a = S::template ld<ROWS>(q + (ptrdiff_t) off * st, ls);
            if (scale) a = S::div(a, dv);
            T sv = a;
            if (fi != 0) {  // uniform across the block
                T b = S::template ld<ROWS>(q + (ptrdiff_t) (off + 1) * st, ls);
                if (scale) b = S::div(b, dv);
                const float fr = __fmul_rn((float) fi, uf);
                sv = S::lerp(fr, a, b);
            }
            acc = S::acc(acc, tap, sv);
        }
    } else if (p >= start && p <= end) {
        // interior pass: the coordinate is carried across taps (imutil.c:803-818)
        // The coordinate chain does not depend on the samples, so the taps are evaluated in
        // batches of 4: coordinates first, then the 8 loads together, then the in-order
        // accumulation (the loads of one voxel would otherwise serialise on L2 latency).
        float c = (float) p;
        int d = -hw;
        for (; d + 3 <= hw; d += 4) {
            float cs[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float step = __fmul_rn((float) (d + j), uf);
                c = __fsub_rn(c, step);
                cs[j] = c;
                c = __fadd_rn(c, step);
            }
            T sv[4];
#pragma unroll
            for (int j = 0; j < 4; j++) sv[j] = samp<T, ROWS>(line, stride, cs[j], n, scale, dv, org, ls);
#pragma unroll
            for (int j = 0; j < 4; j++) acc = S::acc(acc, taps[d + j + hw], sv[j]);
        }
        for (; d <= hw; d++) {
            const float tap = taps[d + hw];
            const float step = __fmul_rn((float) d, uf);
            c = __fsub_rn(c, step);
            acc = S::acc(acc, tap, samp<T, ROWS>(line, stride, c, n, scale, dv, org, ls));
            c = __fadd_rn(c, step);
        }
    } else {
        // boundary pass (imutil.c:833-854)
        for (int d = -hw; d <= hw; d++) {
            const float tap = taps[d + hw];
            const float step = __fmul_rn((float) d, uf);
            float c = __fsub_rn((float) p, step);
            if ((int) c < 0) {
                c = -c;
            } else if ((int) c >= dim_end) {
                c = __fsub_rn(__fsub_rn(__fmul_rn(2.0f, (float) dim_end), c), 0.1f);
            }
            acc = S::acc(acc, tap, samp<T, ROWS>(line, stride, c, n, scale, dv, org, ls));
        }
    }
    return acc;
}

// forward: cross-block max reduction, defined with the max-abs kernels below
__device__ __forceinline__ void block_atomic_max(float m, float *out);

// One axis of the generic 3-pass blur, one thread per voxel. The z pass (EPI) also carries
// the epilogue of a level: max|prev - out| (DoG max-abs) and the 2x decimation.
template <int AXIS, bool EPI>
__global__ void __launch_bounds__(256)
blur_axis_generic_kernel(const float *__restrict__ src, float *__restrict__ dst, int nx, int ny,
                         int nz, const float *__restrict__ taps, int hw, float uf, int uhw,
                         const float *__restrict__ divisor, const float *__restrict__ dog_prev,
                         float *__restrict__ dogmax, float *__restrict__ ds_dst, int sh, int zoff,
                         int zorg) {
    // 3-D launch (x in the block, rows and planes in the grid): no 64-bit div/mod per thread
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int z = blockIdx.z + zoff;
    const size_t idx = ((size_t) z * ny + (size_t) y) * nx + (size_t) x;
    float vmax = 0.0f;
    if (x < nx && y < ny) {
        const int p = AXIS == 0 ? x : (AXIS == 1 ? y : z);
        const int n = AXIS == 0 ? nx : (AXIS == 1 ? ny : nz);
        const size_t stride = AXIS == 0 ? 1 : (AXIS == 1 ? (size_t) nx : (size_t) nx * ny);
        float dv = 0.0f;
        if (divisor != nullptr) dv = *divisor;
        const bool scale = dv != 0.0f;  // im_scale leaves an all-zero image alone (imutil.c:706)
        // z of a z-slab: coordinates of the whole volume, array starting at plane zorg
        const int org = AXIS == 2 ? zorg : 0;
        const float o = conv_point_generic(src + (idx - (size_t) p * stride), stride, p + org,
                                           n + org, taps, hw, uf, uhw, scale, dv, sh, org);
        dst[idx] = o;
        if (EPI) {
            if (dogmax != nullptr) vmax = fabsf(__fsub_rn(dog_prev[idx], o));
            if (ds_dst != nullptr && !((x | y | z) & 1)) {
                const int dnx = nx >> 1, dny = ny >> 1, dnz = nz >> 1;
                const int qx = x >> 1, qy = y >> 1, qz = z >> 1;
                if (qx < dnx && qy < dny && qz < dnz)
                    ds_dst[((size_t) qz * dny + qy) * dnx + qx] = o;
            }
        }
    }
    if (EPI) {
        if (dogmax != nullptr) block_atomic_max(vmax, dogmax);
    }
}

// x pass with the same column of four rows per thread (ny % 4 == 0): the sample coordinates
// depend on x only, so one coordinate chain serves four voxels (scalar loads, coalesced
// across the warp as before).
__global__ void __launch_bounds__(256)
blur_axis_generic_rows4_kernel(const float *__restrict__ src, float *__restrict__ dst, int nx, int ny,
                               int nz, const float *__restrict__ taps, int hw, float uf, int uhw,
                               const float *__restrict__ divisor, int sh, int zoff) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = 4 * (blockIdx.y * blockDim.y + threadIdx.y);
    const int z = blockIdx.z + zoff;
    if (x >= nx || y >= ny) return;
    const size_t row = ((size_t) z * ny + (size_t) y) * nx;
    float dv = 0.0f;
    if (divisor != nullptr) dv = *divisor;
    const bool scale = dv != 0.0f;
    const float4 o = conv_point_generic<float4, true>(src + row, 1, x, nx, taps, hw, uf, uhw, scale, dv, sh, 0,
                                                      (size_t) nx);
    float *q = dst + row + x;
    q[0] = o.x;
    q[(size_t) nx] = o.y;
    q[2 * (size_t) nx] = o.z;
    q[3 * (size_t) nx] = o.w;
}

// y or z pass with four x-adjacent voxels per thread (nx % 4 == 0, 16-byte aligned arrays):
// sample coordinates, fractions and addresses depend on (y, z) only, so they are formed once
// per four voxels and the samples are 16-byte loads (2.5x fewer instructions per voxel than
// the scalar kernel; the arithmetic per voxel is the same sequence, hence the same bits).
template <int AXIS, bool EPI>
__global__ void __launch_bounds__(256)
blur_axis_generic4_kernel(const float *__restrict__ src, float *__restrict__ dst, int nx, int ny,
                          int nz, const float *__restrict__ taps, int hw, float uf, int uhw,
                          const float *__restrict__ dog_prev, float *__restrict__ dogmax,
                          float *__restrict__ ds_dst, int sh, int zoff, int zorg) {
    static_assert(AXIS == 1 || AXIS == 2, "x passes use the scalar kernel");
    const int x = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int z = blockIdx.z + zoff;
    const size_t idx = ((size_t) z * ny + (size_t) y) * nx + (size_t) x;
    float vmax = 0.0f;
    if (x < nx && y < ny) {
        const int p = AXIS == 1 ? y : z;
        const int n = AXIS == 1 ? ny : nz;
        const size_t stride = AXIS == 1 ? (size_t) nx : (size_t) nx * ny;
        const int org = AXIS == 2 ? zorg : 0;
        const float4 o = conv_point_generic<float4>(src + (idx - (size_t) p * stride), stride, p + org,
                                                    n + org, taps, hw, uf, uhw, false, 0.0f, sh, org);
        *reinterpret_cast<float4 *>(dst + idx) = o;
        if (EPI) {
            if (dogmax != nullptr) {
                const float4 q = *reinterpret_cast<const float4 *>(dog_prev + idx);
                vmax = fmaxf(fmaxf(fabsf(__fsub_rn(q.x, o.x)), fabsf(__fsub_rn(q.y, o.y))),
                             fmaxf(fabsf(__fsub_rn(q.z, o.z)), fabsf(__fsub_rn(q.w, o.w))));
            }
            if (ds_dst != nullptr && !((y | z) & 1)) {
                const int dnx = nx >> 1, dny = ny >> 1, dnz = nz >> 1;
                const int qx = x >> 1, qy = y >> 1, qz = z >> 1;
                if (qy < dny && qz < dnz) {
                    float *q = ds_dst + ((size_t) qz * dny + qy) * dnx + qx;
                    if (qx < dnx) q[0] = o.x;
                    if (qx + 1 < dnx) q[1] = o.z;
                }
            }
        }
    }
    if (EPI) {
        if (dogmax != nullptr) block_atomic_max(vmax, dogmax);
    }
}

// im_scale (imutil.c:699-713): v / max|v| as a true division; an all-zero image is left alone.
__global__ void __launch_bounds__(256)
scale_kernel(const float *__restrict__ src, float *__restrict__ dst, size_t n,
             const float *__restrict__ divisor) {
    const float dv = *divisor;
    const bool scale = dv != 0.0f;
    const size_t n4 = n >> 2;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    const float4 *__restrict__ s4 = reinterpret_cast<const float4 *>(src);
    float4 *__restrict__ d4 = reinterpret_cast<float4 *>(dst);
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 v = s4[i];
        if (scale) {
            v.x = __fdiv_rn(v.x, dv); v.y = __fdiv_rn(v.y, dv);
            v.z = __fdiv_rn(v.z, dv); v.w = __fdiv_rn(v.w, dv);
        }
        d4[i] = v;
    }
    for (size_t i = (n4 << 2) + (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        dst[i] = scale ? __fdiv_rn(src[i], dv) : src[i];
}

// ---------------------------------------------------------------------------------------
// 2x decimation (imutil.c:597-613)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
downsample_kernel(const float *__restrict__ src, float *__restrict__ dst, int snx, int sny,
                  int dnx, int dny, int dnz) {
    const size_t total = (size_t) dnx * dny * dnz;
    const size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int x = (int) (idx % dnx);
    const int y = (int) ((idx / dnx) % dny);
    const int z = (int) (idx / ((size_t) dnx * dny));
    dst[idx] = src[((size_t) (2 * z) * sny + (size_t) (2 * y)) * snx + (size_t) (2 * x)];
}

// ---------------------------------------------------------------------------------------
// max |v| and max |a - b| (imutil.c:681-695, sift.c:822-826). Non-negative floats order
// like their bit patterns, so the cross-block reduction is an integer atomicMax.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void block_atomic_max(float m, float *out) {
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    __shared__ float wmax[32];
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;  // 1-D and 2-D blocks
    const int lane = tid & 31, warp = tid >> 5;
    if (lane == 0) wmax[warp] = m;
    __syncthreads();
    if (warp == 0) {
        m = lane < (int) ((blockDim.x * blockDim.y) >> 5) ? wmax[lane] : 0.0f;
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) atomicMax(reinterpret_cast<int *>(out), __float_as_int(m));
    }
}

__global__ void __launch_bounds__(256)
absmax_kernel(const float *__restrict__ a, size_t n, float *__restrict__ out) {
    float m = 0.0f;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    // 16-byte loads, four in flight per thread (a scalar grid-stride loop reached 2.3 TB/s)
    const size_t head = (16 - ((uintptr_t) a & 15)) / 4 % 4;  // scalars before 16-byte alignment
    const size_t h = head < n ? head : n;
    if (tid < h) m = fabsf(a[tid]);
    const float4 *__restrict__ a4 = reinterpret_cast<const float4 *>(a + h);
    const size_t n4 = (n - h) >> 2;
    size_t i = tid;
    for (; i + 3 * stride < n4; i += 4 * stride) {
        const float4 v0 = a4[i], v1 = a4[i + stride], v2 = a4[i + 2 * stride], v3 = a4[i + 3 * stride];
        m = fmaxf(m, fmaxf(fmaxf(fabsf(v0.x), fabsf(v0.y)), fmaxf(fabsf(v0.z), fabsf(v0.w))));
        m = fmaxf(m, fmaxf(fmaxf(fabsf(v1.x), fabsf(v1.y)), fmaxf(fabsf(v1.z), fabsf(v1.w))));
        m = fmaxf(m, fmaxf(fmaxf(fabsf(v2.x), fabsf(v2.y)), fmaxf(fabsf(v2.z), fabsf(v2.w))));
        m = fmaxf(m, fmaxf(fmaxf(fabsf(v3.x), fabsf(v3.y)), fmaxf(fabsf(v3.z), fabsf(v3.w))));
    }
    for (; i < n4; i += stride) {
        const float4 v = a4[i];
        m = fmaxf(m, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
    }
    for (size_t j = h + (n4 << 2) + tid; j < n; j += stride) m = fmaxf(m, fabsf(a[j]));
    block_atomic_max(m, out);
}

__global__ void __launch_bounds__(256)
absdiffmax_kernel(const float *__restrict__ a, const float *__restrict__ b, size_t n,
                  float *__restrict__ out) {
    float m = 0.0f;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        m = fmaxf(m, fabsf(__fsub_rn(a[i], b[i])));
    block_atomic_max(m, out);
}

}  // namespace s3d

#include "blur_zmarch.cuh"
#include "blur_fused_half.cuh"
#include "blur_band.cuh"

namespace s3d {

// ---------------------------------------------------------------------------------------
// Launchers
// ---------------------------------------------------------------------------------------
static inline AxisSpec axis_spec(int n, double unit, int hw) {
    AxisSpec a;
    a.n = n;
    a.uf = (float) (1.0 / unit);                 // imutil.c:754-755 (unit argument is 1.0)
    a.uhw = (int) ceilf((float) hw * a.uf);      // imutil.c:756-757
    // dyadic unit factor 2^-sh on a line short enough that every coordinate is exact
    a.sh = -1;
    for (int k = 0; k <= 8; k++)
        if (a.uf == ldexpf(1.0f, -k) && n < (1 << 20)) a.sh = k;
    return a;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                  const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (!tried) {
        tried = true;
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
                cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn) p;
        else
            cudaGetLastError();
    }
    return fn;
}

// Cut the flattened (column tile, output plane) space of the fused unit-spacing kernel into
// one contiguous range per CTA with equal COST: a chunk (the part of a range inside one
// tile) costs its outputs plus the 2*hw halo planes that warm its z window up (plus the
// hw + 2 prologue planes of a chunk that ends at the top of the volume). Equal z-chunks per
// tile would leave 148 mod tiles SMs idle (128 of 148 CTAs at 256^3). Chunks hold at least
// hw + 1 outputs: a chunk below the top must not reach plane nz-1 with its halo.
static bool fill_chunk_table(ChunkTable &t, int ncta, int tiles, int nzr, int hw, bool top_is_volume_top,
                             int budget, int *max_cost) {
    const int minlen = hw + 1 < nzr ? hw + 1 : nzr;
    const long total = (long) tiles * nzr;
    long pos = 0;
    int worst = 0;
    t.b[0] = 0;
    for (int i = 0; i < ncta; i++) {
        int left = budget, cost = 0;
        while (pos < total) {
            const int rem = nzr - (int) (pos % nzr);  // outputs left in this tile
            const int halo = 2 * hw;
            int take = left - halo;
            if (take < minlen) {
                if (cost > 0) break;
                take = minlen;
            }
            if (take > rem) take = rem;
            if (rem - take > 0 && rem - take < minlen) {  // never leave a stub behind
                if (take - (minlen - (rem - take)) >= minlen) take = rem - minlen;
                else take = rem;
            }
            const int c = take + halo + ((take == rem && top_is_volume_top) ? hw + 2 : 0);
            pos += take;
            left -= c;
            cost += c;
        }
        t.b[i + 1] = (int) pos;
        if (cost > worst) worst = cost;
    }
    *max_cost = worst;
    return pos == total;
}

static const ChunkTable &fused_chunk_table(int ncta, int tiles, int nzr, int hw, bool top) {
    struct Entry { int ncta, tiles, nzr, hw, top; ChunkTable t; };
    static thread_local Entry cache[16];
    static thread_local int used = 0, next = 0;
    for (int i = 0; i < used; i++) {
        const Entry &e = cache[i];
        if (e.ncta == ncta && e.tiles == tiles && e.nzr == nzr && e.hw == hw && e.top == (int) top) return e.t;
    }
    Entry &e = cache[next];
    next = (next + 1) % 16;
    if (used < 16) used++;
    e.ncta = ncta; e.tiles = tiles; e.nzr = nzr; e.hw = hw; e.top = (int) top;
    const long total = (long) tiles * nzr;
    int budget = (int) ((total + ncta - 1) / ncta) + 2 * hw;
    ChunkTable best;
    int best_cost = 0x7fffffff, cost = 0;
    // the first feasible budget is near optimal; a few more are tried because a range may
    // overshoot its budget where a tile's remainder would be shorter than a legal chunk
    for (int tries = 0, feasible = 0; tries < 4096 && feasible < 8; tries++, budget++) {
        ChunkTable t;
        if (!fill_chunk_table(t, ncta, tiles, nzr, hw, top, budget, &cost)) continue;
        feasible++;
        if (cost < best_cost) {
            best_cost = cost;
            best = t;
        }
    }
    if (best_cost == 0x7fffffff) {  // cannot happen (budget = one tile is always feasible)
        for (int i = 0; i <= ncta; i++) best.b[i] = (int) (total * i / ncta / nzr * nzr);
        best.b[ncta] = (int) total;
    }
    for (int i = ncta + 1; i <= kMaxFusedCtas; i++) best.b[i] = best.b[ncta];
    e.t = best;
    return e.t;
}

static thread_local int g_fused_sm_reserve = 0;
void set_fused_sm_reserve(int sms) { g_fused_sm_reserve = sms < 0 ? 0 : (sms > kNumSMs - 1 ? kNumSMs - 1 : sms); }

}  // namespace s3d
extern "C" SIFT3D_CUDA_EXPORT int sift3d_cuda_blur_chunk_table(int ncta, int tiles, int nzr, int hw,
                                                              int top_is_volume_top, int *bounds) {
    using namespace s3d;
    if (ncta < 1 || ncta > kMaxFusedCtas || tiles < 1 || nzr < 1 || hw < 1 || !bounds ||
        (long) tiles * nzr > 0x7fffffffL)
        return -1;
    const ChunkTable &t = fused_chunk_table(ncta, tiles, nzr, hw, top_is_volume_top != 0);
    int worst = 0;
    for (int i = 0; i <= ncta; i++) bounds[i] = t.b[i];
    for (int i = 0; i < ncta; i++) {
        int cost = 0;
        for (long pos = t.b[i]; pos < t.b[i + 1];) {
            const int rem = nzr - (int) (pos % nzr);
            const int take = (int) (t.b[i + 1] - pos < rem ? t.b[i + 1] - pos : rem);
            cost += take + 2 * hw + ((take == rem && top_is_volume_top) ? hw + 2 : 0);
            pos += take;
        }
        if (cost > worst) worst = cost;
    }
    return worst;
}
extern "C" SIFT3D_CUDA_EXPORT int sift3d_cuda_blur_band_table(int n_line, int org, float unit_factor,
                                                             const sift3d_cuda_filter *f, float *w,
                                                             size_t w_capacity, int *base, int *band) {
    using namespace s3d;
    if (!f || !w || !base || !band || f->width < 1 || f->width > SIFT3D_CUDA_MAX_TAPS || !(f->width & 1) ||
        org < 0 || n_line - org < 2 || !(unit_factor > 0.0f))
        return -1;
    BandKey K;
    memset(&K, 0, sizeof(K));
    K.n_line = n_line;
    K.org = org;
    K.hw = f->width / 2;
    K.uf = unit_factor;
    for (int i = 0; i < f->width; i++) K.taps[i] = f->taps[i];
    std::vector<float> wT;
    std::vector<int> b;
    int B = 0;
    if (build_band(K, wT, b, B)) return 1;
    *band = B;
    if (wT.size() > w_capacity) return 1;
    memcpy(w, wT.data(), wT.size() * sizeof(float));
    memcpy(base, b.data(), b.size() * sizeof(int));
    return 0;
}
namespace s3d {

template <int HW, int R, int NT, bool EXACT>
static int launch_fused_t(cudaStream_t st, const FusedParams &P) {
    using C = FusedCfg<HW>;
    using Z = ZCfg<HW, R, NT>;
    using L = typename Z::L;
    constexpr int TY = Z::TY;
    constexpr int TROWS = L::TROWS;
    constexpr size_t SMEM = L::SMEM;
    if ((size_t) P.nx * P.ny * P.nz >= ((size_t) 1 << 31)) return 1;  // 32-bit voxel offsets in the kernel
    if (P.ny < TY) return 1;
    EncodeTiledFn enc = encode_tiled_fn();
    if (enc == nullptr) return 1;
    CUtensorMap tmap;
    const cuuint64_t gdim[3] = {(cuuint64_t) P.nx, (cuuint64_t) P.ny, (cuuint64_t) P.nz};
    const cuuint64_t gstr[2] = {(cuuint64_t) P.nx * 4, (cuuint64_t) P.nx * P.ny * 4};
    const cuuint32_t box[3] = {(cuuint32_t) C::BW, (cuuint32_t) TROWS, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void *) P.src, gdim, gstr, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return 1;
    static PerDeviceOnce attr_once;
    if (attr_once.run([&]() -> int {
            S3D_CK(cudaFuncSetAttribute(blur_zmarch_kernel<HW, R, NT, EXACT>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int) SMEM));
            return 0;
        }))
        return -1;
    const int tiles = ((P.nx + kFTX - 1) / kFTX) * ((P.ny + TY - 1) / TY);
    const int nzr = P.z1 - P.z0;
    // one wave of CTAs, but no CTA with less than ~8 outputs of one tile
    long ncta = (long) (kNumSMs - g_fused_sm_reserve) * Z::OCC;
    const long most = (long) tiles * (nzr / (HW + 1 > 8 ? HW + 1 : 8) > 0 ? nzr / (HW + 1 > 8 ? HW + 1 : 8) : 1);
    if (ncta > most) ncta = most;
    if (ncta > kMaxFusedCtas) ncta = kMaxFusedCtas;
    if ((long) tiles * nzr > 0x7fffffffL) return 1;
    const ChunkTable &tab = fused_chunk_table((int) ncta, tiles, nzr, HW, P.z1 == P.nz);
    blur_zmarch_kernel<HW, R, NT, EXACT><<<(unsigned) ncta, NT, SMEM, st>>>(tmap, P, tab);
    S3D_CK(cudaGetLastError());
    return 0;
}

// Thread / tile shape of the z-marching kernel per half-width:
//   variant 0   one row of four voxels per thread, 256 threads (32 x 32 tiles)
//   variant 1   two rows, 256 threads (32 x 64 tiles)
//   variant 2   four rows, 128 threads (32 x 64 tiles; half-widths <= 3)
//   variant 3   two rows, 128 threads (32 x 32 tiles)
// More rows per thread amortise the per-plane fixed work; the running z sums (8 * R * HW
// registers) bound R for the wide filters. $SIFT3D_CUDA_ZV = one digit per half-width 1..8.
static int zmarch_variant(int hw, int ny) {
    static const char *env = getenv("SIFT3D_CUDA_ZV");
    int v = hw <= 4 ? 1 : (hw <= 6 ? 3 : 1);
    if (env != nullptr && (int) strlen(env) >= hw && env[hw - 1] >= '0' && env[hw - 1] <= '3') v = env[hw - 1] - '0';
    if (v == 2 && hw > 3) v = 1;
    if ((v == 1 || v == 2) && ny < 64) v = 0;
    return v;
}

template <int HW, bool EXACT>
static int launch_fused_v(cudaStream_t st, const FusedParams &P, int variant) {
    if (variant == 0) return launch_fused_t<HW, 1, 256, EXACT>(st, P);
    if (variant == 3) return launch_fused_t<HW, 2, 128, EXACT>(st, P);
    if constexpr (HW <= 3) {
        if (variant == 2) return launch_fused_t<HW, 4, 128, EXACT>(st, P);
    }
    return launch_fused_t<HW, 2, 256, EXACT>(st, P);
}

// Returns 0 if launched, 1 if this blur is not eligible for the fused kernel, -1 on error.
template <bool EXACT>
static int launch_fused(cudaStream_t st, FusedParams &P, int hw) {
    P.zchunk = P.z1 - P.z0;
    const int v = zmarch_variant(hw, P.ny);
    switch (hw) {
    case 1: return launch_fused_v<1, EXACT>(st, P, v);
    case 2: return launch_fused_v<2, EXACT>(st, P, v);
    case 3: return launch_fused_v<3, EXACT>(st, P, v);
    case 4: return launch_fused_v<4, EXACT>(st, P, v);
    case 5: return launch_fused_v<5, EXACT>(st, P, v);
    case 6: return launch_fused_v<6, EXACT>(st, P, v);
    case 7: return launch_fused_v<7, EXACT>(st, P, v);
    case 8: return launch_fused_v<8, EXACT>(st, P, v);
    default: return 1;
    }
}

template <int HW, int TY>
static int launch_half_t(cudaStream_t st, const FusedParams &P, int nzc) {
    using C = HalfCfg<HW>;
    constexpr int TROWS = TY + 2 * C::YH;
    constexpr int TILE_FLOATS = TROWS * C::BW;
    constexpr int TILE_STRIDE = (TILE_FLOATS + 31) / 32 * 32;
    constexpr int NUR = 2 * TY + 2 * HW - 1;
    constexpr size_t SMEM = (size_t) (kNB * TILE_STRIDE + 2 * TROWS * kFTX + 2 * NUR * kFTX) * 4 +
                            kNB * 8 + (HW + 1) * 4 + 16;
    EncodeTiledFn enc = encode_tiled_fn();
    if (enc == nullptr) return 1;
    CUtensorMap tmap;
    const cuuint64_t gdim[3] = {(cuuint64_t) P.nx, (cuuint64_t) P.ny, (cuuint64_t) P.nz};
    const cuuint64_t gstr[2] = {(cuuint64_t) P.nx * 4, (cuuint64_t) P.nx * P.ny * 4};
    const cuuint32_t box[3] = {(cuuint32_t) C::BW, (cuuint32_t) TROWS, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void *) P.src, gdim, gstr, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return 1;
    static PerDeviceOnce attr_once;
    if (attr_once.run([&]() -> int {
            S3D_CK(cudaFuncSetAttribute(blur_fused_half_kernel<HW, TY>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int) SMEM));
            return 0;
        }))
        return -1;
    dim3 grid((unsigned) ((P.nx + kFTX - 1) / kFTX), (unsigned) ((P.ny + TY - 1) / TY), (unsigned) nzc);
    blur_fused_half_kernel<HW, TY><<<grid, 8 * TY, SMEM, st>>>(tmap, P);
    S3D_CK(cudaGetLastError());
    return 0;
}

// Spacing-1/2 fused kernel; same return convention and chunking rules as launch_fused, with
// the z footprint uhw = ceil(hw / 2).
template <int TY>
static int launch_half(cudaStream_t st, FusedParams &P, int hw, int occ) {
    const int uhw = (hw + 1) / 2;
    const int tiles = ((P.nx + kFTX - 1) / kFTX) * ((P.ny + TY - 1) / TY);
    int nzc = (kNumSMs * occ) / tiles;
    if (nzc < 1) nzc = 1;
    const int nzr = P.z1 - P.z0;  // planes to produce
    int zc = (nzr + nzc - 1) / nzc;
    if (zc < uhw + 1) zc = uhw + 1;
    nzc = (nzr + zc - 1) / zc;
    while (nzc > 1 && nzr - (nzc - 1) * zc < uhw + 1) {
        zc++;
        nzc = (nzr + zc - 1) / zc;
    }
    P.zchunk = zc;
    switch (hw) {
    case 1: return launch_half_t<1, TY>(st, P, nzc);
    case 2: return launch_half_t<2, TY>(st, P, nzc);
    case 3: return launch_half_t<3, TY>(st, P, nzc);
    case 4: return launch_half_t<4, TY>(st, P, nzc);
    case 5: return launch_half_t<5, TY>(st, P, nzc);
    case 6: return launch_half_t<6, TY>(st, P, nzc);
    case 7: return launch_half_t<7, TY>(st, P, nzc);
    case 8: return launch_half_t<8, TY>(st, P, nzc);
    default: return 1;
    }
}

int launch_blur(cudaStream_t st, const float *src, float *dst, float *tmp_a, float *tmp_b,
                int nx, int ny, int nz, const double units[3], const sift3d_cuda_filter &f,
                const float *d_taps, const float *d_divisor, float *d_dogmax_slot,
                const float *dog_prev, float *ds_dst, bool force_generic, uint64_t *launches,
                int z0, int z1, int zorg, bool exact, int vol_nz, bool no_alloc) {
    const int hw = f.width / 2;
    if (z1 < 0) {  // whole volume
        z0 = 0;
        z1 = nz;
    }
    const AxisSpec ax = axis_spec(nx, units[0], hw);
    const AxisSpec ay = axis_spec(ny, units[1], hw);
    const AxisSpec az = axis_spec(nz + zorg, units[2], hw);
    // z-slab mode: a CUT face must keep the filter footprint inside the array (the planes beyond
    // the owned range are halo filled by the neighbouring device); a face of the array that is a
    // face of the volume (zorg == 0 below, zorg + nz == vol_nz above) takes the boundary rules.
    const bool bottom_is_volume = zorg == 0;
    const bool top_is_volume = vol_nz <= 0 ? false : zorg + nz == vol_nz;
    const int need = az.uhw + 2;
    bool fused_ok = true;  // the fused kernels apply the top boundary rule in the chunk that ends at nz only
    if (z0 < 0 || z1 > nz || z0 >= z1 || (z0 > 0 && !bottom_is_volume && z0 < need) ||
        (z1 < nz && nz - z1 < need && !top_is_volume) ||
        (vol_nz <= 0 && ((z0 > 0 && z0 < 2 * hw + 2) || (z1 < nz && nz - z1 < 2 * hw + 2)))) {
        S3D_ERR("sift3d_cuda: z range [%d, %d) of %d planes leaves no room for a %d-tap halo\n", z0, z1,
                nz, f.width);
        return -1;
    }
    if (z1 < nz && nz - z1 < need) fused_ok = false;
    const size_t total = (size_t) nx * ny * nz;
    const unsigned blocks = (unsigned) ((total + 255) / 256);

    // Fused single-launch path: unit tap spacing on every axis, TMA-compatible rows, taps
    // symmetric bit for bit (the kernel shares the products of twin taps).
    // Tile height: 32-row tiles at 3 CTAs x 256 threads per SM (<= 85 registers, hw <= 4) are
    // ~5 % faster than one 64-row CTA of 512 threads; the wider filters need the registers.
    static const int ty_env = getenv("SIFT3D_CUDA_TY") ? atoi(getenv("SIFT3D_CUDA_TY")) : 0;
    const int ty_half = (ty_env == 32 || ny < 64) ? 32 : 64;
    bool symmetric = true;
    for (int i = 0; i < hw; i++) symmetric = symmetric && f.taps[i] == f.taps[f.width - 1 - i];
    if (!force_generic && fused_ok && symmetric && ax.uf == 1.0f && ay.uf == 1.0f && az.uf == 1.0f && hw >= 1 &&
        hw <= 8 && f.width <= kMaxParamTaps && nx % 4 == 0 && nx >= kFTX && ny >= 32 &&
        ny >= 2 * hw + 2 && nz >= 2 * hw + 2 && (((uintptr_t) src | (uintptr_t) dst | (uintptr_t) ds_dst) & 15) == 0 &&
        (d_dogmax_slot == nullptr || dog_prev == src) &&
        (d_divisor == nullptr || (tmp_a != nullptr && tmp_a != src))) {
        FusedParams P;
        P.src = src;
        P.dst = dst;
        P.ds_dst = ds_dst;
        P.dogmax = d_dogmax_slot;
        P.out_scale = nullptr;
        if (d_divisor != nullptr && exact) {
            // im_scale (imutil.c:699-713) as its own streaming pass: inside the blur every
            // voxel would be divided once per register window that holds it (3x for this
            // filter) and the reciprocal's latency stalls the tile pipeline (145 us against
            // 85 us for the next, wider, filter); one pass costs ~25 us.
            scale_kernel<<<kNumSMs * 8, 256, 0, st>>>(src, tmp_a, total, d_divisor);
            if (launches) *launches += 1;
            P.src = tmp_a;
        } else if (d_divisor != nullptr) {
            // fast mode: the blur is linear, so v / max is applied to the OUTPUT as one multiply
            // by 1 / max (one more rounding, ~6e-8 relative) and the streaming pass disappears
            P.out_scale = d_divisor;
        }
        P.nx = nx; P.ny = ny; P.nz = nz;
        P.z0 = z0; P.z1 = z1; P.zorg = zorg;
        P.zchunk = nz;
        for (int i = 0; i < f.width; i++) P.taps.k[i] = f.taps[i];
        for (int i = f.width; i < kMaxParamTaps; i++) P.taps.k[i] = 0.0f;
        if (d_dogmax_slot != nullptr) S3D_CK(cudaMemsetAsync(d_dogmax_slot, 0, sizeof(float), st));
        const int r = exact ? launch_fused<true>(st, P, hw) : launch_fused<false>(st, P, hw);
        if (r < 0) return -1;
        if (r == 0) {
            if (launches) *launches += 1;
            return 0;
        }
        // not taken after all: the generic path below divides on load from the original
    }

    // Fused single-launch path for tap spacing 1/2 on every axis (octave 1): exact mode. In fast mode
    // the three banded passes below take octave 1 as well -- its levels (8 MB at 256^3) stay in L2
    // between the passes and the banded form is far fewer instructions than the upsampled-line
    // construction of this kernel: 256^3 step 1.97 -> 1.90 ms per volume ($SIFT3D_CUDA_HALF_FUSED=1:
    // the fused kernel in fast mode too).
    static const bool half_fused = getenv("SIFT3D_CUDA_HALF_FUSED") != nullptr;
    if ((exact || half_fused) && !force_generic && fused_ok && symmetric && ax.uf == 0.5f && ay.uf == 0.5f && az.uf == 0.5f && hw >= 1 &&
        hw <= 8 && d_divisor == nullptr && nx % 4 == 0 && nx >= kFTX && ny >= ty_half && nx < (1 << 20) &&
        ny < (1 << 20) && nz < (1 << 20) && ny >= hw + 3 && nz >= hw + 3 &&
        ((uintptr_t) src & 15) == 0 && (d_dogmax_slot == nullptr || dog_prev == src)) {
        FusedParams P;
        P.src = src;
        P.dst = dst;
        P.ds_dst = ds_dst;
        P.dogmax = d_dogmax_slot;
        P.out_scale = nullptr;
        P.nx = nx; P.ny = ny; P.nz = nz;
        P.z0 = z0; P.z1 = z1; P.zorg = zorg;
        P.zchunk = nz;
        for (int i = 0; i < f.width; i++) P.taps.k[i] = f.taps[i];
        for (int i = f.width; i < kMaxParamTaps; i++) P.taps.k[i] = 0.0f;
        if (d_dogmax_slot != nullptr) S3D_CK(cudaMemsetAsync(d_dogmax_slot, 0, sizeof(float), st));
        const int r = ty_half == 64 ? launch_half<64>(st, P, hw, 1) : launch_half<32>(st, P, hw, 2);
        if (r < 0) return -1;
        if (r == 0) {
            if (launches) *launches += 1;
            return 0;
        }
    }

    const bool want_max = d_dogmax_slot != nullptr && dog_prev != nullptr;
    if (want_max) S3D_CK(cudaMemsetAsync(d_dogmax_slot, 0, sizeof(float), st));

    // Fast mode, any tap spacing: the three passes as banded multiply-adds (blur_band.cuh).
    static const bool no_band = getenv("SIFT3D_CUDA_NO_BAND") != nullptr;
    if (!exact && !force_generic && !no_band && !no_alloc && nx % 4 == 0 && ny % 4 == 0 && nx >= 16 &&
        (((uintptr_t) src | (uintptr_t) tmp_a | (uintptr_t) tmp_b | (uintptr_t) dst | (uintptr_t) dog_prev) & 15) == 0 &&
        tmp_a != nullptr && tmp_b != nullptr) {
        BandAxis bx_, by_, bz_;
        int r = get_band(nx, 0, ax.uf, f, 4, 4, bx_);
        if (r == 0) r = get_band(ny, 0, ay.uf, f, 1, kBandR, by_);
        if (r == 0) r = get_band(nz + zorg, zorg, az.uf, f, 1, kBandR, bz_);
        if (r < 0) return -1;
        if (r == 0) {
            // x and y passes cover the planes the z blocks of [z0, z1) read
            const int kz0 = z0 / kBandR, kz1 = (z1 + kBandR - 1) / kBandR;
            const int reach = bz_.BB + kBandR;
            const int zlo = z0 - reach > 0 ? z0 - reach : 0;
            const int zhi = z1 + reach < nz ? z1 + reach : nz;
            const int nbx = nx / 4;  // blocks of four x (x pass), float4 columns (y, z passes)
            int tx = 8;
            while (tx < nbx && tx < 32) tx <<= 1;
            {
                const int rows = ny % 8 == 0 ? 8 : 4;
                unsigned ty = 128 / tx;
                if (ty > (unsigned) (ny / rows)) ty = (unsigned) (ny / rows);
                const dim3 block((unsigned) tx, ty, 1);
                const dim3 grid((unsigned) ((nbx + tx - 1) / tx), (unsigned) ((ny / rows + ty - 1) / ty),
                                (unsigned) (zhi - zlo));
                if (rows == 8) band_x_kernel<8><<<grid, block, 0, st>>>(src, tmp_a, nx, ny, bx_, d_divisor, zlo);
                else band_x_kernel<4><<<grid, block, 0, st>>>(src, tmp_a, nx, ny, bx_, d_divisor, zlo);
            }
            int vx = 8;
            while (vx < nbx && vx < 64) vx <<= 1;
            const dim3 vblock((unsigned) vx, (unsigned) (256 / vx), 1);
            const dim3 grid_y((unsigned) ((nbx + vx - 1) / vx), (unsigned) ((by_.nblk + vblock.y - 1) / vblock.y),
                              (unsigned) (zhi - zlo));
            const dim3 grid_z(grid_y.x, (unsigned) ((ny + vblock.y - 1) / vblock.y), (unsigned) (kz1 - kz0));
            const size_t ring_bytes = (size_t) kBandDepth * 256 * sizeof(float4);
            band_blk_kernel<1, false><<<grid_y, vblock, ring_bytes, st>>>(tmp_a, tmp_b, nx, ny, nz, by_, nullptr, nullptr,
                                                                nullptr, zlo, 0, ny);
            band_blk_kernel<2, true><<<grid_z, vblock, ring_bytes, st>>>(tmp_b, dst, nx, ny, nz, bz_, dog_prev,
                                                               want_max ? d_dogmax_slot : nullptr, ds_dst, kz0,
                                                               z0, z1);
            if (launches) *launches += 3;
            S3D_CK(cudaGetLastError());
            return 0;
        }
    }
    int bx = 32;
    while (bx < nx && bx < 256) bx <<= 1;
    const dim3 gblock((unsigned) bx, (unsigned) (256 / bx), 1);
    // x and y passes cover the z range plus the z footprint of the last pass
    const int zlo = z0 - az.uhw - 1 > 0 ? z0 - az.uhw - 1 : 0;
    const int zhi = z1 + az.uhw + 1 < nz ? z1 + az.uhw + 1 : nz;
    const dim3 ggrid_xy((unsigned) ((nx + bx - 1) / bx), (unsigned) ((ny + gblock.y - 1) / gblock.y),
                        (unsigned) (zhi - zlo));
    const dim3 ggrid_z(ggrid_xy.x, ggrid_xy.y, (unsigned) (z1 - z0));
    // im_scale as its own streaming pass into dst (free until the z pass writes it): on load
    // every sample of the x pass would be divided, 2 * (2 hw + 1) divisions per voxel
    const float *xsrc = src;
    if (d_divisor != nullptr && dst != src && dst != tmp_a && dst != tmp_b && z0 == 0 && z1 == nz) {
        scale_kernel<<<kNumSMs * 8, 256, 0, st>>>(src, dst, total, d_divisor);
        if (launches) *launches += 1;
        xsrc = dst;
        d_divisor = nullptr;
    }
    static const bool no_vec = getenv("SIFT3D_CUDA_NO_VEC4") != nullptr;
    if (!no_vec && ny % 4 == 0) {
        const unsigned by4 = gblock.y < (unsigned) (ny / 4) ? gblock.y : (unsigned) (ny / 4);
        const dim3 rblock((unsigned) bx, by4 > 0 ? by4 : 1, 1);
        const dim3 rgrid((unsigned) ((nx + bx - 1) / bx), (unsigned) ((ny / 4 + rblock.y - 1) / rblock.y),
                         (unsigned) (zhi - zlo));
        blur_axis_generic_rows4_kernel<<<rgrid, rblock, 0, st>>>(xsrc, tmp_a, nx, ny, nz, d_taps, hw, ax.uf,
                                                             ax.uhw, d_divisor, ax.sh, zlo);
    } else {
        blur_axis_generic_kernel<0, false><<<ggrid_xy, gblock, 0, st>>>(
            xsrc, tmp_a, nx, ny, nz, d_taps, hw, ax.uf, ax.uhw, d_divisor, nullptr, nullptr, nullptr, ax.sh, zlo, 0);
    }
    const bool vec4 = !no_vec && nx % 4 == 0 && nx >= 16 &&
                      (((uintptr_t) tmp_a | (uintptr_t) tmp_b | (uintptr_t) dst | (uintptr_t) dog_prev |
                        (uintptr_t) 0) & 15) == 0;
    if (vec4) {
        int bx4 = 8;
        while (bx4 < nx / 4 && bx4 < 64) bx4 <<= 1;
        const dim3 vblock((unsigned) bx4, (unsigned) (256 / bx4), 1);
        const dim3 vgrid_xy((unsigned) ((nx / 4 + bx4 - 1) / bx4), (unsigned) ((ny + vblock.y - 1) / vblock.y),
                            (unsigned) (zhi - zlo));
        const dim3 vgrid_z(vgrid_xy.x, vgrid_xy.y, (unsigned) (z1 - z0));
        blur_axis_generic4_kernel<1, false><<<vgrid_xy, vblock, 0, st>>>(
            tmp_a, tmp_b, nx, ny, nz, d_taps, hw, ay.uf, ay.uhw, nullptr, nullptr, nullptr, ay.sh, zlo, 0);
        blur_axis_generic4_kernel<2, true><<<vgrid_z, vblock, 0, st>>>(
            tmp_b, dst, nx, ny, nz, d_taps, hw, az.uf, az.uhw, dog_prev,
            want_max ? d_dogmax_slot : nullptr, ds_dst, az.sh, z0, zorg);
    } else {
        blur_axis_generic_kernel<1, false><<<ggrid_xy, gblock, 0, st>>>(
            tmp_a, tmp_b, nx, ny, nz, d_taps, hw, ay.uf, ay.uhw, nullptr, nullptr, nullptr, nullptr, ay.sh, zlo, 0);
        blur_axis_generic_kernel<2, true><<<ggrid_z, gblock, 0, st>>>(
            tmp_b, dst, nx, ny, nz, d_taps, hw, az.uf, az.uhw, nullptr, dog_prev,
            want_max ? d_dogmax_slot : nullptr, ds_dst, az.sh, z0, zorg);
    }
    if (launches) *launches += 3;
    S3D_CK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------
// The last octaves of a pyramid are a few thousand voxels each, yet every level costs three
// launches and the octaves form a serial chain (octave o+1 starts from level 2 of octave o):
// at 256^3 the octaves of 16^3 and 8^3 voxels alone took 320 us of launch latency at the end
// of a 1.7 ms detection. This kernel runs a whole chain of such octaves in ONE single-CTA
// launch: levels live in shared memory, each blur is the same x / y / z sequence of
// conv_point_generic calls as the generic kernels (so the bits are the same), with the DoG
// max-abs and the 2x decimation into the next octave done in place. All levels are also
// written to global memory for the extrema / orientation / descriptor kernels.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTinyThreads)
tiny_octaves_kernel(const TinyChainParams P) {
    extern __shared__ __align__(16) float tiny_smem[];
    __shared__ float red[kTinyThreads / 32];
    const int tid = threadIdx.x;
    float *cur = tiny_smem, *nxt = tiny_smem + kTinyMaxVox, *ta = tiny_smem + 2 * kTinyMaxVox,
          *tb = tiny_smem + 3 * kTinyMaxVox;
    float *const seed = tiny_smem + 4 * kTinyMaxVox;  // level -1 of the next octave
    for (int t = 0; t < P.noct; t++) {
        const TinyOctave &O = P.oct[t];
        const int nx = O.nx, ny = O.ny, nz = O.nz;
        const int n = nx * ny * nz;
        // rows of the (y, z) list are nx floats: 16-byte aligned, and groups of four rows stay
        // inside the volume, when nx and ny * nz are multiples of 4
        const bool vec4 = nx % 4 == 0 && (ny * nz) % 4 == 0;
        // level -1: from global for the first octave, else already decimated into `cur`
        if (t == 0)
            for (int i = tid; i < n; i += kTinyThreads) cur[i] = O.lvl[0][i];
        __syncthreads();
        for (int s = 0; s < P.nfilt; s++) {
            const float *__restrict__ taps = P.taps[s];
            const int hw = P.hw[s];
            for (int axis = 0; axis < 3; axis++) {
                const float *__restrict__ src = axis == 0 ? cur : (axis == 1 ? ta : tb);
                float *__restrict__ dst = axis == 0 ? ta : (axis == 1 ? tb : nxt);
                const float uf = O.uf[axis];
                const int uhw = (int) ceilf((float) hw * uf);  // imutil.c:756-757
                const int sh = O.sh[axis];
                const int len = axis == 0 ? nx : (axis == 1 ? ny : nz);
                const size_t stride = axis == 0 ? 1 : (axis == 1 ? (size_t) nx : (size_t) nx * ny);
                if (P.banded) {
                    // fast mode: at these octaves the taps are a fraction of a voxel apart (2^-4 and
                    // below), so a 17-tap filter reads 3 - 5 samples: the banded form is a handful of
                    // multiply-adds per voxel instead of 17 interpolated taps
                    // (the chain is a sequence of ~30 dependent passes of ~4 items per thread: what
                    // counts is the latency of one item, so its loads are issued together)
                    const TinyBand B = P.band[t][s][axis];
                    if (nx % 4 == 0) {
                        // item = four x-adjacent voxels. The whole chain runs on ONE SM (1024 threads,
                        // ~40 dependent phases), so a phase costs its instruction count: per four
                        // voxels one index computation, BB weight loads and BB 16-byte samples
                        // (y, z) or BB scalar samples against 16-byte weights (x).
                        const float4 *__restrict__ w4 = reinterpret_cast<const float4 *>(B.w);
                        for (int i4 = tid; i4 < n / 4; i4 += kTinyThreads) {
                            const int i = 4 * i4;
                            const int yz = i / nx, x = i - yz * nx;
                            const int z = yz / ny, y = yz - z * ny;
                            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (axis == 0) {
                                const int k = x >> 2;
                                const float *__restrict__ q = src + (size_t) (i - x) + (size_t) B.bb[k];
                                const float4 *__restrict__ wk = w4 + k;
#pragma unroll 4
                                for (int r = 0; r < B.BB; r++) {
                                    const float4 wr = __ldg(wk + (size_t) r * B.nblk);
                                    const float v = q[r];
                                    acc.x = fmaf(wr.x, v, acc.x); acc.y = fmaf(wr.y, v, acc.y);
                                    acc.z = fmaf(wr.z, v, acc.z); acc.w = fmaf(wr.w, v, acc.w);
                                }
                            } else {
                                const int p = axis == 1 ? y : z;
                                const int k = p >> 2;
                                const float *__restrict__ q = src + ((size_t) i - (size_t) p * stride) + (size_t) B.bb[k] * stride;
                                const float *__restrict__ wq = B.w + (size_t) k * 4 + (p & 3);
#pragma unroll 4
                                for (int r = 0; r < B.BB; r++) {
                                    const float wv = __ldg(wq + (size_t) r * B.nblk * 4);
                                    const float4 v = *reinterpret_cast<const float4 *>(q + (size_t) r * stride);
                                    acc.x = fmaf(wv, v.x, acc.x); acc.y = fmaf(wv, v.y, acc.y);
                                    acc.z = fmaf(wv, v.z, acc.z); acc.w = fmaf(wv, v.w, acc.w);
                                }
                            }
                            *reinterpret_cast<float4 *>(dst + i) = acc;
                        }
                    } else {
                        for (int i = tid; i < n; i += kTinyThreads) {
                            const int yz = i / nx, x = i - yz * nx;
                            const int z = yz / ny, y = yz - z * ny;
                            const int p = axis == 0 ? x : (axis == 1 ? y : z);
                            const int k = p >> 2, ii = p & 3;
                            const float *__restrict__ q = src + ((size_t) i - (size_t) p * stride) + (size_t) B.bb[k] * stride;
                            const float *__restrict__ wq = B.w + (size_t) k * 4 + ii;
                            float acc = 0.0f;
#pragma unroll 4
                            for (int r = 0; r < B.BB; r++)
                                acc = fmaf(__ldg(wq + (size_t) r * B.nblk * 4), q[(size_t) r * stride], acc);
                            dst[i] = acc;
                        }
                    }
                } else if (vec4 && axis > 0) {
                    // four x-adjacent voxels per item: one coordinate chain, 16-byte accesses
                    for (int i4 = tid; i4 < n / 4; i4 += kTinyThreads) {
                        const int i = 4 * i4;
                        const int y = (i / nx) % ny, z = i / (nx * ny);
                        const int p = axis == 1 ? y : z;
                        *reinterpret_cast<float4 *>(dst + i) = conv_point_generic<float4>(
                            src + ((size_t) i - (size_t) p * stride), stride, p, len, taps, hw, uf, uhw, false,
                            0.0f, sh);
                    }
                } else if (vec4) {
                    // x pass: the same column of four rows per item
                    for (int i4 = tid; i4 < n / 4; i4 += kTinyThreads) {
                        const int x = i4 % nx, r4 = i4 / nx;  // rows 4*r4 .. 4*r4+3 of the (y, z) row list
                        const size_t row = (size_t) (4 * r4) * nx;
                        const float4 o = conv_point_generic<float4, true>(src + row, 1, x, len, taps, hw, uf, uhw,
                                                                          false, 0.0f, sh, 0, (size_t) nx);
                        float *q = dst + row + x;
                        q[0] = o.x; q[nx] = o.y; q[2 * nx] = o.z; q[3 * nx] = o.w;
                    }
                } else {
                    for (int i = tid; i < n; i += kTinyThreads) {
                        const int x = i % nx, y = (i / nx) % ny, z = i / (nx * ny);
                        const int p = axis == 0 ? x : (axis == 1 ? y : z);
                        dst[i] = conv_point_generic(src + ((size_t) i - (size_t) p * stride), stride, p, len,
                                                    taps, hw, uf, uhw, false, 0.0f, sh);
                    }
                }
                __syncthreads();
            }
            // level s is in `nxt`: global copy, DoG max-abs of level s-1, decimation
            float *__restrict__ g = O.lvl[s + 1];
            float vmax = 0.0f;
            const bool want_max = O.dogmax[s] != nullptr;
            if (n % 4 == 0) {
                for (int i = 4 * tid; i < n; i += 4 * kTinyThreads) {
                    const float4 o = *reinterpret_cast<const float4 *>(nxt + i);
                    *reinterpret_cast<float4 *>(g + i) = o;
                    if (want_max) {
                        const float4 c = *reinterpret_cast<const float4 *>(cur + i);
                        vmax = fmaxf(fmaxf(vmax, fmaxf(fabsf(__fsub_rn(c.x, o.x)), fabsf(__fsub_rn(c.y, o.y)))),
                                     fmaxf(fabsf(__fsub_rn(c.z, o.z)), fabsf(__fsub_rn(c.w, o.w))));
                    }
                }
            } else {
                for (int i = tid; i < n; i += kTinyThreads) {
                    const float o = nxt[i];
                    g[i] = o;
                    if (want_max) vmax = fmaxf(vmax, fabsf(__fsub_rn(cur[i], o)));
                }
            }
            if (want_max) {
                for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
                if ((tid & 31) == 0) red[tid >> 5] = vmax;
                __syncthreads();
                if (tid == 0) {
                    float m = 0.0f;
                    for (int w = 0; w < kTinyThreads / 32; w++) m = fmaxf(m, red[w]);
                    *O.dogmax[s] = m;
                }
            }
            if (s == P.ds_level && t + 1 < P.noct) {
                // im_downsample_2x (imutil.c:597-613) into the next octave's level -1
                const TinyOctave &N = P.oct[t + 1];
                float *__restrict__ gn = N.lvl[0];
                const int dn = N.nx * N.ny * N.nz;
                for (int i = tid; i < dn; i += kTinyThreads) {
                    const int x = i % N.nx, y = (i / N.nx) % N.ny, z = i / (N.nx * N.ny);
                    const float v = nxt[((size_t) (2 * z) * ny + (size_t) (2 * y)) * nx + (size_t) (2 * x)];
                    gn[i] = v;
                    seed[i] = v;  // becomes `cur` of the next octave
                }
            }
            __syncthreads();
            float *sw = cur; cur = nxt; nxt = sw;
        }
        if (t + 1 < P.noct) {
            const TinyOctave &N = P.oct[t + 1];
            const int dn = N.nx * N.ny * N.nz;
            __syncthreads();
            for (int i = tid; i < dn; i += kTinyThreads) cur[i] = seed[i];
        }
    }
}

int launch_tiny_octaves(cudaStream_t st, const TinyChainParams &P, uint64_t *launches) {
    const size_t dyn = (size_t) (4 * kTinyMaxVox + kTinyMaxVox / 8 + 64) * sizeof(float);
    static PerDeviceOnce attr_once;
    if (attr_once.run([&]() -> int {
            S3D_CK(cudaFuncSetAttribute(tiny_octaves_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int) dyn));
            return 0;
        }))
        return -1;
    tiny_octaves_kernel<<<1, kTinyThreads, dyn, st>>>(P);
    if (launches) *launches += 1;
    S3D_CK(cudaGetLastError());
    return 0;
}

int tiny_band(int n, double unit, const sift3d_cuda_filter &f, TinyBand *out) {
    const AxisSpec a = axis_spec(n, unit, f.width / 2);
    BandAxis A;
    const int r = get_band(n, 0, a.uf, f, 1, 4, A);
    if (r) return r;
    out->w = reinterpret_cast<const float *>(A.w4);
    out->bb = A.bb;
    out->BB = A.BB;
    out->nblk = A.nblk;
    return 0;
}

// Host: sampling parameters of one axis for the tiny chain (same as the generic launcher)
void tiny_axis(int n, double unit, float *uf, int *sh) {
    const AxisSpec a = axis_spec(n, unit, 1);
    *uf = a.uf;
    *sh = a.sh;
}

int launch_downsample(cudaStream_t st, const float *src, float *dst, int snx, int sny, int snz,
                      uint64_t *launches) {
    const int dnx = snx / 2, dny = sny / 2, dnz = snz / 2;
    const size_t total = (size_t) dnx * dny * dnz;
    downsample_kernel<<<(unsigned) ((total + 255) / 256), 256, 0, st>>>(src, dst, snx, sny, dnx,
                                                                       dny, dnz);
    if (launches) *launches += 1;
    S3D_CK(cudaGetLastError());
    return 0;
}

int launch_absmax(cudaStream_t st, const float *src, size_t n, float *d_out, uint64_t *launches) {
    S3D_CK(cudaMemsetAsync(d_out, 0, sizeof(float), st));
    size_t blocks = (n + 256 * 8 - 1) / (256 * 8);
    if (blocks > (size_t) kNumSMs * 8) blocks = (size_t) kNumSMs * 8;
    if (blocks < 1) blocks = 1;
    absmax_kernel<<<(unsigned) blocks, 256, 0, st>>>(src, n, d_out);
    if (launches) *launches += 1;
    S3D_CK(cudaGetLastError());
    return 0;
}

int launch_absdiffmax(cudaStream_t st, const float *a, const float *b, size_t n, float *d_out,
                      uint64_t *launches) {
    S3D_CK(cudaMemsetAsync(d_out, 0, sizeof(float), st));
    size_t blocks = (n + 256 * 8 - 1) / (256 * 8);
    if (blocks > (size_t) kNumSMs * 8) blocks = (size_t) kNumSMs * 8;
    if (blocks < 1) blocks = 1;
    absdiffmax_kernel<<<(unsigned) blocks, 256, 0, st>>>(a, b, n, d_out);
    if (launches) *launches += 1;
    S3D_CK(cudaGetLastError());
    return 0;
}

}  // namespace s3d

// ---------------------------------------------------------------------------------------
// One blur without any allocation (sift3d_cuda_blur_static, include/sift3d_cuda.h). It has
// no counterpart in the reference; it was added for
// callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops,
// harnesses that forbid it). The scratch volumes of the generic 3-pass blur and the device
// copy of the taps live in a file-scope __device__ array that every GPU loading this library
// holds until exit. In host code the array's NAME is only a host-side stand-in: its device
// address is taken once per device with cudaGetSymbolAddress (that device current) and it is
// this address that goes to the kernels.
// ---------------------------------------------------------------------------------------
#include <mutex>

namespace s3d {

constexpr size_t kStaticBlurMaxVox = (size_t) SIFT3D_CUDA_STATIC_BLUR_MAX_VOXELS;
constexpr size_t kStaticTapsFloats = 256;  // >= SIFT3D_CUDA_MAX_TAPS, keeps the volumes 256-byte aligned
static_assert(kStaticTapsFloats >= SIFT3D_CUDA_MAX_TAPS && kStaticTapsFloats % 64 == 0, "taps slot");
// [taps | scratch volume A | scratch volume B]
static __device__ __align__(256) float g_static_blur_scratch[kStaticTapsFloats + 2 * kStaticBlurMaxVox];

__global__ void static_blur_taps_kernel(const sift3d_cuda_filter f, float *__restrict__ dst) {
    const int i = threadIdx.x;
    if (i < f.width) dst[i] = f.taps[i];
}

// Device address of the scratch array on `device` (which must be current), or nullptr.
static float *static_blur_scratch(int device) {
    static std::mutex mu;
    static float *ptr[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    if (ptr[device] == nullptr) {
        float *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **) &p, g_static_blur_scratch);
        if (e != cudaSuccess || p == nullptr) {
            S3D_ERR("sift3d_cuda_blur_static: cudaGetSymbolAddress failed on device %d: %s\n", device,
                    cudaGetErrorString(e));
            cudaGetLastError();
            return nullptr;
        }
        ptr[device] = p;
    }
    return ptr[device];
}

}  // namespace s3d

extern "C" SIFT3D_CUDA_EXPORT int sift3d_cuda_blur_static(int device, const float *d_src, float *d_dst,
                                                         int nx, int ny, int nz, const double units[3],
                                                         const sift3d_cuda_filter *f, int exact,
                                                         void *stream, int sync) {
    using namespace s3d;
    if (!d_src || !d_dst || !units || !f || nx < 2 || ny < 2 || nz < 2 || f->width < 1 ||
        f->width > SIFT3D_CUDA_MAX_TAPS || !(f->width & 1) || d_src == d_dst) {
        S3D_ERR("sift3d_cuda_blur_static: invalid argument\n");
        return -1;
    }
    const size_t n = (size_t) nx * ny * nz;
    if (n > kStaticBlurMaxVox) {
        S3D_ERR("sift3d_cuda_blur_static: %d x %d x %d = %zu voxels exceed the static scratch of %zu voxels "
                "(SIFT3D_CUDA_STATIC_BLUR_MAX_VOXELS); use sift3d_cuda_blur or a detector\n",
                nx, ny, nz, n, kStaticBlurMaxVox);
        return -1;
    }
    S3D_CK(cudaSetDevice(device));
    float *base = static_blur_scratch(device);
    if (base == nullptr) return -1;
    float *d_taps = base;
    float *tmp_a = base + kStaticTapsFloats;
    float *tmp_b = tmp_a + kStaticBlurMaxVox;
    cudaStream_t st = (cudaStream_t) stream;
    // Nothing in the scratch has to start at zero: the taps slot and every scratch voxel the
    // later passes read are rewritten by this call before they are read.
    static_blur_taps_kernel<<<1, SIFT3D_CUDA_MAX_TAPS, 0, st>>>(*f, d_taps);
    S3D_CK(cudaGetLastError());
    if (launch_blur(st, d_src, d_dst, tmp_a, tmp_b, nx, ny, nz, units, *f, d_taps, nullptr, nullptr,
                    nullptr, nullptr, false, nullptr, 0, -1, 0, exact != 0, 0, /* no_alloc: the band tables are
                    allocated at first use, so the literal kernels take the general spacings */ true))
        return -1;
    if (sync) S3D_CK(cudaStreamSynchronize(st));
    return 0;
}
