// blur_zmarch.cuh -- the roofline kernel (v5): one launch produces one Gaussian level from the
// previous one (x, y and z passes fused; 1 HBM read + 1 HBM write per voxel), for levels whose
// tap spacing is exactly one voxel on every axis (octave 0 with units (1,1,1)).
//
// Reference semantics: apply_Sep_FIR_filter imutil.c:1127-1206 -> convolve_sep_gen
// imutil.c:742-861 (x pass, then y, then z; fp32 between passes; taps accumulated in ascending
// d; mirror at the left face, mirror "- 0.1" at the right face, imutil.c:846-849). With unit
// spacing every sample has frac == 0 and all three boundary rules reduce to a plain FIR over
// an EXTENDED line: E[-i] = E[i] and E[dim_end + m] = lerp(E[dim_end-m-1], E[dim_end-m], fr_m),
// m = 0..HW (interior outputs never touch index dim_end, so the virtual value is consistent).
//
// Two arithmetic modes (template EXACT):
//   EXACT  separate fp32 multiply and add roundings in the reference's order: 0 ULP against
//          the reference build (x86-64 without FMA);
//   fast   one FFMA per tap (one rounding instead of two, ~1e-7 relative): inside the 1e-5
//          voxel tolerance of the parity contract, half the FP32 instructions. im_scale
//          (imutil.c:699-713) is then folded into the first blur as one multiply per output
//          by 1 / max|image| (the blur is linear), which removes a whole streaming pass.
//
// Structure:
//   * a CTA owns a 32 x TY column tile and marches DOWN z through its range of the flattened
//     (tile, plane) space (ChunkTable, cost balanced on the host, one wave of persistent CTAs);
//   * plane tiles with x/y halo are staged by TMA (cp.async.bulk.tensor.3d + mbarrier, ring of
//     4, issued by thread 0; TMA zero-fills outside the volume). For tiles whose x halo reaches
//     a volume face the last warps (which own no second x item) rewrite the halo columns of
//     the staged tile in shared memory into the extended line, one plane ahead of the x pass
//     -- so the x/y/z passes are pure FIR code with no boundary logic at all;
//   * x pass: item = (tile row, 4-column group): LDS.128 -> registers -> STS.128 into a
//     double-buffered x-filtered tile. The y faces are handled by remapping the source row of
//     an item (mirror) or blending two x-filtered rows (virtual rows), offsets precomputed per
//     chunk. Every tile produces ALL its 32 x TY outputs correctly (tiles shifted inwards at
//     the high faces overlap and store identical values), so there are no store masks;
//   * y pass: thread (x = lane, 4 consecutive rows) gathers 4 + 2*HW rows (LDS.32, conflict
//     free); z pass in registers, scatter form (descending march: the running sums receive
//     their taps in ascending d, as convolve_sep_gen accumulates them);
//   * epilogue: max|G(s-1) - G(s)| (sift.c:822-826) and the 2x decimation into the next octave
//     (imutil.c:591-617) fused into the store;
//   * the x pass of plane n+1 and the y/z pass of plane n run between the same pair of
//     barriers: one __syncthreads per plane.
#pragma once

#include <cuda.h>

#include "common.cuh"

namespace s3d {

constexpr int kFTX = 32;  // tile width (x): one warp lane per column in the y/z passes
constexpr int kNB = 4;    // TMA ring (plane tiles in flight)

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t) __cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *tmap, int x, int y, int z,
                                            uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst)),
        "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
        : "memory");
}
// generic-proxy writes to a staged tile (halo fix-up) before the async proxy reuses the buffer
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

struct FusedParams {
    const float *src;        // previous level (also the minuend of the DoG max)
    float *dst;              // this level
    float *ds_dst;           // next octave's level -1 (2x decimation of dst) or nullptr
    float *dogmax;           // slot for max|src - dst| or nullptr
    const float *out_scale;  // fast mode, first blur: *out_scale = max|image|; outputs are
                             // multiplied by 1 / *out_scale (im_scale); nullptr: no scaling
    int nx, ny, nz;
    int z0, z1;              // output planes [z0, z1)
    int zorg;                // z-slab: plane of the whole volume that array plane 0 is (else 0)
    int zchunk;              // (spacing-1/2 kernel) outputs per z-chunk
    TapsParam taps;          // taps in the parameter bank
};

template <int HW>
struct FusedCfg {
    static constexpr int NW = 2 * HW + 1;           // taps
    static constexpr int HWP = (HW + 3) / 4 * 4;    // x halo rounded up to 4: TMA boxes whose start / width
                                                    // are not 16-byte multiples fault ("illegal instruction")
    static constexpr int BW = kFTX + 2 * HWP;       // staged tile width
    static constexpr int NIN = 4 + 2 * HWP;         // inputs a 4-wide x item reads
    static constexpr int NINY = 4 + 2 * HW;         // rows a 4-high y item reads
};

// Shared-memory layout of the z-marching kernel (kernel and launcher agree through this)
template <int HW, int TY>
struct ZmarchLayout {
    static constexpr int TROWS = TY + 2 * HW;
    static constexpr int TILE_FLOATS = TROWS * FusedCfg<HW>::BW;
    static constexpr int TILE_STRIDE = (TILE_FLOATS + 31) / 32 * 32;
    static constexpr int XF_STRIDE = TROWS * kFTX;
    static constexpr size_t SMEM = (size_t) (kNB * TILE_STRIDE + 2 * XF_STRIDE) * 4 + kNB * 8 +
                                   2 * (HW + 1) * 4 + 16;
};

__device__ __forceinline__ float lerp1(const float a, const float b, const float fr) {
    return __fadd_rn(__fmul_rn(__fsub_rn(1.0f, fr), a), __fmul_rn(fr, b));
}

// Fraction of the right-boundary mirror for a tap landing m samples past dim_end:
// c = dim_end + m -> 2*dim_end - c - 0.1f, frac = c - (int) c (imutil.c:846-853).
__device__ __forceinline__ float right_frac(int dim_end, int m) {
    const float c = (float) (dim_end + m);
    const float cm = __fsub_rn(__fsub_rn(__fmul_rn(2.0f, (float) dim_end), c), 0.1f);
    return __fsub_rn(cm, (float) (int) cm);
}

// tap index with the symmetric twin folded onto the lower half (the Gaussian is symmetric bit
// for bit, checked on the host), so k[t] and k[2*HW - t] are the same constant-bank operand
template <int HW>
__device__ __forceinline__ constexpr int fold(int t) {
    return t <= HW ? t : 2 * HW - t;
}

// one tap: acc + k * v (EXACT: two roundings as the reference; fast: one FFMA)
template <bool EXACT>
__device__ __forceinline__ float tap_acc(const float acc, const float k, const float v) {
    if (EXACT) return __fadd_rn(acc, __fmul_rn(k, v));
    return __fmaf_rn(k, v, acc);
}
// first tap of a sum: EXACT keeps the reference's 0.0f + k*v (it turns -0 into +0)
template <bool EXACT>
__device__ __forceinline__ float tap_first(const float k, const float v) {
    if (EXACT) return __fadd_rn(0.0f, __fmul_rn(k, v));
    return __fmul_rn(k, v);
}

// Shared-memory accesses on 32-bit shared addresses with immediate offsets: address arithmetic
// stays one 32-bit add per plane instead of a generic -> shared conversion per access group.
// (volatile: a staged buffer is rewritten every other plane, the loads must not be merged)
template <int OFF>
__device__ __forceinline__ float4 lds128(const uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "r"(a), "n"(OFF));
    return v;
}
template <int OFF>
__device__ __forceinline__ float lds32(const uint32_t a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(a), "n"(OFF));
    return v;
}
__device__ __forceinline__ void sts128(const uint32_t a, const float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// x pass of one tile row for one 4-column group: 4 x-filtered outputs. `p` is the shared
// address of tile column 4*g of the row (tile column 0 <-> x0 - HWP); the row already is the
// extended line.
template <int HW, bool EXACT, int Q = 0>
__device__ __forceinline__ void xrow4_load(const uint32_t p, float (&in)[FusedCfg<HW>::NIN]) {
    if constexpr (Q < FusedCfg<HW>::NIN / 4) {
        const float4 t4 = lds128<16 * Q>(p);
        in[4 * Q] = t4.x; in[4 * Q + 1] = t4.y; in[4 * Q + 2] = t4.z; in[4 * Q + 3] = t4.w;
        xrow4_load<HW, EXACT, Q + 1>(p, in);
    }
}
template <int HW, bool EXACT>
__device__ __forceinline__ float4 xrow4_fir(const float (&in)[FusedCfg<HW>::NIN], const TapsParam &k) {
    using C = FusedCfg<HW>;
    constexpr int NW = C::NW, HWP = C::HWP;
    float o[4];
#pragma unroll
    for (int c = 0; c < 4; c++) {
        float a = tap_first<EXACT>(k.k[0], in[c + HWP + HW]);  // tap d = t - HW samples x - d
#pragma unroll
        for (int t = 1; t < NW; t++) a = tap_acc<EXACT>(a, k.k[fold<HW>(t)], in[c + HWP + HW - t]);
        o[c] = a;
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}
template <int HW, bool EXACT>
__device__ __forceinline__ float4 xrow4(const uint32_t p, const TapsParam &k) {
    float in[FusedCfg<HW>::NIN];
    xrow4_load<HW, EXACT>(p, in);
    return xrow4_fir<HW, EXACT>(in, k);
}

// z pass, scatter form: feed one y-filtered value; returns the output that is now complete.
// A[t] holds the sum of taps 0..t of the output that is t+1 taps old.
template <int HW, bool EXACT>
__device__ __forceinline__ float zfeed(float (&A)[2 * HW], const float v, const TapsParam &k) {
    if (EXACT) {
        float P[HW + 1];
#pragma unroll
        for (int j = 0; j <= HW; j++) P[j] = __fmul_rn(k.k[j], v);
        const float out = __fadd_rn(A[2 * HW - 1], P[0]);
#pragma unroll
        for (int t = 2 * HW - 1; t >= 1; t--) A[t] = __fadd_rn(A[t - 1], P[fold<HW>(t)]);
        A[0] = __fadd_rn(0.0f, P[0]);
        return out;
    } else {
        const float out = __fmaf_rn(k.k[0], v, A[2 * HW - 1]);
#pragma unroll
        for (int t = 2 * HW - 1; t >= 1; t--) A[t] = __fmaf_rn(k.k[fold<HW>(t)], v, A[t - 1]);
        A[0] = __fmul_rn(k.k[0], v);
        return out;
    }
}

// Chunk table: CTA i produces the outputs [b[i], b[i+1]) of the flattened (column tile,
// output plane) space, tile-major: unit u is plane z0 + u % nzr of tile u / nzr. The host
// cuts the space so that every CTA has the same cost (outputs + 2*HW halo planes per chunk,
// a chunk being the part of a CTA's range inside one tile), see fused_chunk_table().
constexpr int kMaxFusedCtas = 3 * kNumSMs;
struct ChunkTable {
    int b[kMaxFusedCtas + 1];
};

// mbarrier / TMA helpers on precomputed 32-bit shared addresses (the generic -> shared
// conversion reads SR_CgaCtaId; done once per kernel, not once per plane)
__device__ __forceinline__ uint32_t opaque_u32(uint32_t v) {
    asm volatile("mov.u32 %0, %0;" : "+r"(v));
    return v;
}
__device__ __forceinline__ void mbar_expect_tx_a(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void tma_load_3d_a(uint32_t dst, const CUtensorMap *tmap, int x, int y, int z,
                                              uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
        "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(bar)
        : "memory");
}

// Configuration of the z-marching kernel: R rows of 4 x-adjacent voxels per thread in the y / z
// passes, NT threads (8 column groups x NT/8 row groups), hence 32 x TY tiles, TY = R * NT / 8.
template <int HW, int R, int NT>
struct ZCfg {
    static constexpr int TY = R * NT / 8;
    using L = ZmarchLayout<HW, TY>;
    // CTAs per SM: by shared memory (227 KB usable) and by a register estimate (the running
    // z sums, 4R * 2HW, plus ~64 for the x / y passes and addresses)
    static constexpr int kRegs = 8 * R * HW + 64;
    static constexpr int kOccS = (int) ((size_t) 227 * 1024 / (L::SMEM + 1024));
    static constexpr int kOccR = 65536 / (NT * (kRegs > 255 ? 255 : kRegs));
    static constexpr int kOcc0 = kOccS < kOccR ? kOccS : kOccR;
    static constexpr int OCC = kOcc0 < 1 ? 1 : (kOcc0 > 4 ? 4 : kOcc0);
};

// y pass, streaming form: rows J = R-1+2HW .. 0 of the thread's 4 columns of the x-filtered
// tile, descending, so that every output receives its taps in ascending order (tap t of output
// c is row c + 2HW - t)
template <int HW, int R, bool EXACT, int J>
__device__ __forceinline__ void yrows(const uint32_t p, float (&a)[R][4], const TapsParam &k) {
    if constexpr (J >= 0) {
        const float4 v = lds128<J * kFTX * 4>(p);
#pragma unroll
        for (int c = 0; c < R; c++) {
            const int t = c + 2 * HW - J;
            if (t == 0) {
                a[c][0] = tap_first<EXACT>(k.k[0], v.x);
                a[c][1] = tap_first<EXACT>(k.k[0], v.y);
                a[c][2] = tap_first<EXACT>(k.k[0], v.z);
                a[c][3] = tap_first<EXACT>(k.k[0], v.w);
            } else if (t > 0 && t <= 2 * HW) {
                const float kt = k.k[fold<HW>(t)];
                a[c][0] = tap_acc<EXACT>(a[c][0], kt, v.x);
                a[c][1] = tap_acc<EXACT>(a[c][1], kt, v.y);
                a[c][2] = tap_acc<EXACT>(a[c][2], kt, v.z);
                a[c][3] = tap_acc<EXACT>(a[c][3], kt, v.w);
            }
        }
        yrows<HW, R, EXACT, J - 1>(p, a, k);
    }
}

// y pass in two phases. Loads: rows J = 0 .. R-1+2HW of the thread's 4 columns.
template <int HW, int R, int J = 0>
__device__ __forceinline__ void yload(const uint32_t p, float4 (&in)[R + 2 * HW]) {
    if constexpr (J < R + 2 * HW) {
        in[J] = lds128<J * kFTX * 4>(p);
        yload<HW, R, J + 1>(p, in);
    }
}
// FIR: every output receives its taps in ascending order (tap t of output c is row c + 2HW - t)
template <int HW, int R, bool EXACT>
__device__ __forceinline__ void yfir(const float4 (&in)[R + 2 * HW], float (&a)[R][4], const TapsParam &k) {
#pragma unroll
    for (int c = 0; c < R; c++) {
        const float4 v0 = in[c + 2 * HW];
        a[c][0] = tap_first<EXACT>(k.k[0], v0.x);
        a[c][1] = tap_first<EXACT>(k.k[0], v0.y);
        a[c][2] = tap_first<EXACT>(k.k[0], v0.z);
        a[c][3] = tap_first<EXACT>(k.k[0], v0.w);
    }
#pragma unroll
    for (int t = 1; t <= 2 * HW; t++) {
        const float kt = k.k[fold<HW>(t)];
#pragma unroll
        for (int c = 0; c < R; c++) {
            const float4 v = in[c + 2 * HW - t];
            a[c][0] = tap_acc<EXACT>(a[c][0], kt, v.x);
            a[c][1] = tap_acc<EXACT>(a[c][1], kt, v.y);
            a[c][2] = tap_acc<EXACT>(a[c][2], kt, v.z);
            a[c][3] = tap_acc<EXACT>(a[c][3], kt, v.w);
        }
    }
}

// Fused blur, unit tap spacing on all three axes. Thread (g = tid & 7, rg = tid >> 3) owns
// the 4 columns x0 + 4g .. 4g+3 of the rows y0 + R*rg .. R*rg + R-1 in the y and z passes:
// every shared-memory and global access of those passes is a 16-byte one.
template <int HW, int R, int NT, bool EXACT>
__global__ void __launch_bounds__(NT, ZCfg<HW, R, NT>::OCC)
blur_zmarch_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ FusedParams P,
                   const __grid_constant__ ChunkTable tab) {
    using C = FusedCfg<HW>;
    using Z = ZCfg<HW, R, NT>;
    using L = typename Z::L;
    constexpr int TY = Z::TY;
    constexpr int HWP = C::HWP, BW = C::BW, NIN = C::NIN;
    constexpr int TROWS = L::TROWS;               // staged rows (y halo)
    constexpr uint32_t TILE_BYTES = L::TILE_FLOATS * 4;
    constexpr int TILE_STRIDE = L::TILE_STRIDE;   // floats, 128-byte aligned TMA destinations
    constexpr int XF_STRIDE = L::XF_STRIDE;       // floats between the two x-filtered tiles
    constexpr int RPI = NT / 8;                   // tile rows covered by one round of x items
    constexpr int kItems = (TROWS + RPI - 1) / RPI;
    constexpr int NV = 4 * R;                     // voxels per thread
    // Loads in flight: the shared-memory loads of a pass are issued before the arithmetic of the
    // pass in front of it where the registers allow (a warp otherwise sits out one LDS latency
    // per x item and one per y pass, with 4 warps per scheduler to cover it):
#ifndef S3D_XMODE
#define S3D_XMODE 0
#endif
#ifndef S3D_YFIRST
#define S3D_YFIRST 0
#endif
    // x pass: 0 = inputs of an item right before its arithmetic, 1 = inputs of item k+1 before
    // the arithmetic of item k, 2 = inputs of all items first
    constexpr int XMODE = (S3D_XMODE == 2 && kItems * NIN > 36) ? 1 : S3D_XMODE;
    constexpr bool XALL = XMODE == 2;
    constexpr bool YFIRST = S3D_YFIRST && (R + 2 * HW) * 4 <= 40;  // y-pass inputs before the x pass

    extern __shared__ __align__(128) unsigned char smem_raw[];
    float *tile = reinterpret_cast<float *>(smem_raw);                 // [kNB][TROWS][BW]
    float *xf = tile + kNB * TILE_STRIDE;                              // [2][TROWS][32]
    uint64_t *mbar = reinterpret_cast<uint64_t *>(xf + 2 * XF_STRIDE); // [kNB]
    float *frx = reinterpret_cast<float *>(mbar + kNB);                // [HW + 1]
    float *fry = frx + (HW + 1);                                       // [HW + 1]
    __shared__ float wmax_sh[NT / 32];

    const int tid = threadIdx.x;
    const int g = tid & 7, rg = tid >> 3;
    const int nx = P.nx, ny = P.ny, nz = P.nz;
    const int dex = nx - 1, dey = ny - 1, dez = nz - 1;
    const int nzr = P.z1 - P.z0;
    const int ntx = (nx + kFTX - 1) / kFTX;
    const unsigned plane_stride = (unsigned) nx * (unsigned) ny;  // (the launcher keeps nvox < 2^31)
    const int dnx = nx >> 1, dny = ny >> 1, dnz = nz >> 1;
    const uint32_t mbar0 = opaque_u32(smem_u32(mbar));
    const uint32_t tile0 = opaque_u32(smem_u32(tile));
    const uint32_t xf0 = tile0 + (uint32_t) (kNB * TILE_STRIDE * 4);

    if (tid == 0) {
#pragma unroll
        for (int b = 0; b < kNB; b++) mbar_init(&mbar[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid <= HW) {
        frx[tid] = right_frac(dex, tid);
        fry[tid] = right_frac(dey, tid);
    }
    __syncthreads();

    float oscale = 1.0f;  // im_scale folded into the output (fast mode, first blur)
    if (!EXACT && P.out_scale != nullptr) {
        const float dv = *P.out_scale;
        if (dv != 0.0f) oscale = __fdiv_rn(1.0f, dv);  // an all-zero image is left alone (imutil.c:706)
    }

    // Stream elements (plane tiles) are numbered through the whole kernel: element E lives in
    // ring slot E & 3 and completes phase (E >> 2) & 1 of that slot's mbarrier.
    uint32_t ebase = 0;
    float vmax = 0.0f;

    int pos = tab.b[blockIdx.x];
    const int pend = tab.b[blockIdx.x + 1];
    while (pos < pend) {
        const int tl = pos / nzr;
        const int zoff = pos - tl * nzr;
        const int zlen = min(pend - pos, nzr - zoff);
        pos += zlen;
        const int tyi = tl / ntx, txi = tl - tyi * ntx;
        // tiles are shifted inwards at the high faces so every tile is full
        const int x0 = min(txi * kFTX, nx - kFTX);
        const int y0 = min(tyi * TY, ny - TY);
        const int zb = P.z0 + zoff;  // outputs [zb, ze) of this tile
        const int ze = zb + zlen;
        const bool top = ze == nz;
        // the x halo of this tile reaches a volume face: the staged tiles are rewritten into the extended line
        const bool xfix = x0 < HW || x0 + kFTX - 1 + HW >= dex;

        // Stream of fed positions q (z_out = q + HW). Top chunk: a prologue walks planes
        // dez-HW-1 .. dez upwards to build the virtual planes dez+HW .. dez (fed in that
        // order), then the main stream feeds real planes dez-1 downwards. Other chunks:
        // planes ze-1+HW downwards. Negative positions reflect (plane -q).
        const int npro = top ? HW + 2 : 0;
        const int qstart = top ? dez - 1 : ze - 1 + HW;
        const int N = npro + (qstart - (zb - HW) + 1);

        // ---- TMA issue (thread 0) and x-face fix-up (one tile row per thread)
        auto issue = [&](int n) {  // stage stream element n
            const uint32_t s = (ebase + (uint32_t) n) & (kNB - 1);
            int pl;
            if (n < npro) {
                pl = dez - HW - 1 + n;
            } else {
                const int q = qstart - (n - npro);
                pl = q < 0 ? -q : q;
            }
            fence_proxy_async();
            mbar_expect_tx_a(mbar0 + 8 * s, TILE_BYTES);
            tma_load_3d_a(tile0 + s * (TILE_STRIDE * 4), &tmap, x0 - HWP, y0 - HW, pl, mbar0 + 8 * s);
        };
        // Extended line in x of element n: tile column c <-> x = x0 - HWP + c. The rows of the
        // staged tile are rewritten in place, one plane ahead of the x pass.
        auto fix = [&](int n) {
            const uint32_t e = ebase + (uint32_t) n;
            if (tid < TROWS) mbar_wait_a(mbar0 + 8 * (e & (kNB - 1)), (e >> 2) & 1u);
            for (int fr_ = tid; fr_ < TROWS; fr_ += NT) {
                float *row = tile + (e & (kNB - 1)) * TILE_STRIDE + fr_ * BW;
                if (x0 < HW) {  // E[-i] = E[i]
                    for (int c = HWP - HW; c < HWP - x0; c++) row[c] = row[2 * (HWP - x0) - c];
                }
                if (x0 + kFTX - 1 + HW >= dex) {
                    // E[dim_end + m] = lerp(E[dim_end-m-1], E[dim_end-m], fr_m); m = 0 last: it
                    // replaces E[dim_end] itself, the source of no other m
                    const int cd = dex - x0 + HWP;  // tile column of x = dim_end
                    for (int m = HW; m >= 0; m--)
                        if (cd + m <= HWP + kFTX - 1 + HW)
                            row[cd + m] = lerp1(row[cd - m - 1], row[cd - m], frx[m]);
                }
            }
            fence_proxy_async();  // these generic writes precede the TMA refill of the buffer
        };
        if (tid == 0) {
            for (int n = 0; n < kNB && n < N; n++) issue(n);
        }

        float A[NV][2 * HW];
#pragma unroll
        for (int c = 0; c < NV; c++)
#pragma unroll
            for (int t = 0; t < 2 * HW; t++) A[c][t] = 0.0f;

        // ---- x-pass items of this thread (chunk invariant): item k is tile row r = rg + k*RPI,
        // column group g. Source row(s) of tile row r, global row gy = y0 - HW + r:
        //   gy < 0        mirror: the row of -gy
        //   gy >= dey     virtual row dey + m = lerp(row dey-m-1, row dey-m, fry[m]) of the
        //                 x-FILTERED rows (the y pass of the reference reads x-filtered data)
        uint32_t srcA[kItems];  // byte offset of the (first) source row inside a staged tile
#pragma unroll
        for (int k = 0; k < kItems; k++) {
            const int r = rg + k * RPI;
            const int gy = y0 - HW + r;
            int ra = r;
            if (gy < 0) ra = r - 2 * gy;
            else if (gy >= dey) ra = r - 2 * (gy - dey);
            srcA[k] = (uint32_t) (ra * BW + 4 * g) * 4u;
        }
        const uint32_t xdst = xf0 + (uint32_t) (rg * kFTX + 4 * g) * 4u;
        // the tile holds virtual rows (uniform: it touches the high y face)
        const bool yvirt = y0 + TY - 1 + HW >= dey;
        const bool last_item = rg + (kItems - 1) * RPI < TROWS;

        const int x = x0 + 4 * g;
        const int ybase = y0 + R * rg;
        // decimation targets: even z, even rows, even x (components .x and .z)
        unsigned dmask = 0;
        if (P.ds_dst != nullptr) {
#pragma unroll
            for (int c = 0; c < R; c++)
                if (!((ybase + c) & 1) && ((ybase + c) >> 1) < dny) dmask |= 1u << c;
        }
        // element (ze - 1, ybase, x): the first output stored; one plane down per stored plane
        unsigned goff = (unsigned) (ze - 1) * plane_stride + (unsigned) ybase * (unsigned) nx + (unsigned) x;
        const bool want_max = P.dogmax != nullptr;
        const uint32_t ysrc = xf0 + (uint32_t) ((R * rg) * kFTX + 4 * g) * 4u;

        // x pass of stream element n into xf[n & 1]
        auto xpass = [&](const int n) {
            const uint32_t e = ebase + (uint32_t) n;
            const uint32_t s = e & (kNB - 1);
            const uint32_t tb = tile0 + s * (uint32_t) (TILE_STRIDE * 4);
            const uint32_t xd = xdst + ((n & 1) ? (uint32_t) (XF_STRIDE * 4) : 0u);
            mbar_wait_a(mbar0 + 8 * s, (e >> 2) & 1u);
            float xin[XALL ? kItems : 2][NIN];
            if (XALL) {
#pragma unroll
                for (int k = 0; k < kItems; k++)
                    if (k < kItems - 1 || last_item) xrow4_load<HW, EXACT>(tb + srcA[k], xin[k]);
            } else if (XMODE == 1) {
                xrow4_load<HW, EXACT>(tb + srcA[0], xin[0]);
            }
#pragma unroll
            for (int k = 0; k < kItems; k++) {
                if (k == kItems - 1 && !last_item) break;
                if (XMODE == 1 && k + 1 < kItems && (k + 1 < kItems - 1 || last_item))
                    xrow4_load<HW, EXACT>(tb + srcA[k + 1], xin[(k + 1) & 1]);
                if (XMODE == 0) xrow4_load<HW, EXACT>(tb + srcA[k], xin[k & 1]);
                float4 o = xrow4_fir<HW, EXACT>(xin[XALL ? k : (k & 1)], P.taps);
                if (yvirt) {
                    const int m = y0 - HW + rg + k * RPI - dey;
                    if (m >= 0) {
                        const float4 a = xrow4<HW, EXACT>(tb + srcA[k] - (uint32_t) (BW * 4), P.taps);
                        const float fr = fry[m];
                        o = make_float4(lerp1(a.x, o.x, fr), lerp1(a.y, o.y, fr), lerp1(a.z, o.z, fr),
                                        lerp1(a.w, o.w, fr));
                    }
                }
                sts128(xd + (uint32_t) (k * RPI * kFTX * 4), o);
            }
        };
        // one stream element: x pass of element n+1 (if any), y pass of element n
        auto plane = [&](const int n, const bool more, float (&yv)[R][4]) {
            const uint32_t ya = ysrc + ((n & 1) ? (uint32_t) (XF_STRIDE * 4) : 0u);
            if constexpr (YFIRST) {
                float4 yin[R + 2 * HW];
                yload<HW, R>(ya, yin);
                if (more) xpass(n + 1);
                yfir<HW, R, EXACT>(yin, yv, P.taps);
            } else {
                if (more) xpass(n + 1);
                yrows<HW, R, EXACT, R - 1 + 2 * HW>(ya, yv, P.taps);
            }
        };

        if (xfix) {
            fix(0);
            if (1 < N) fix(1);
            __syncthreads();
        }
        xpass(0);
        __syncthreads();
        if (tid == 0 && kNB < N) issue(kNB);

        int n = 0;
        if (top) {
            // ---- prologue (top chunk): y-filtered planes dez-HW-1 .. dez, ascending; from the
            // second one on, the virtual plane dez + m = lerp(plane dez-m-1, plane dez-m) is fed
            float prev[R][4];
#pragma unroll
            for (int c = 0; c < R; c++)
#pragma unroll
                for (int i = 0; i < 4; i++) prev[c][i] = 0.0f;
            for (; n < npro; n++) {
                float yv[R][4];
                plane(n, true, yv);  // (the main stream is never empty)
                if (n > 0) {
                    const float fr = right_frac(dez + P.zorg, HW + 1 - n);
#pragma unroll
                    for (int c = 0; c < R; c++)
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const float v = lerp1(prev[c][i], yv[c][i], fr);
                            (void) zfeed<HW, EXACT>(A[4 * c + i], v, P.taps);  // outputs beyond the volume
                        }
                }
#pragma unroll
                for (int c = 0; c < R; c++)
#pragma unroll
                    for (int i = 0; i < 4; i++) prev[c][i] = yv[c][i];
                if (xfix && n + 2 < N) fix(n + 2);
                __syncthreads();
                // the buffer of element n+1 was consumed by the x pass of this iteration
                if (tid == 0 && n + 1 + kNB < N) issue(n + 1 + kNB);
            }
        }
        // ---- main stream: the first fed planes only warm the window up (their outputs lie
        // at or above ze); from then on every plane completes the output one plane lower
        int zo = qstart + HW;  // output completed by element n
        for (; n < N; n++, zo--) {
            const bool store = zo < ze;
            // source voxels of the DoG max: requested a whole plane step before they are used
            float4 pv[R];
            if (want_max && store) {
#pragma unroll
                for (int c = 0; c < R; c++)
                    pv[c] = __ldg(reinterpret_cast<const float4 *>(P.src + (goff + (unsigned) (c * nx))));
            }
            float yv[R][4];
            plane(n, n + 1 < N, yv);
            float o[R][4];
#pragma unroll
            for (int c = 0; c < R; c++)
#pragma unroll
                for (int i = 0; i < 4; i++) o[c][i] = zfeed<HW, EXACT>(A[4 * c + i], yv[c][i], P.taps);
            if (store) {
                if (!EXACT && P.out_scale != nullptr) {
#pragma unroll
                    for (int c = 0; c < R; c++)
#pragma unroll
                        for (int i = 0; i < 4; i++) o[c][i] = __fmul_rn(o[c][i], oscale);
                }
#pragma unroll
                for (int c = 0; c < R; c++)
                    *reinterpret_cast<float4 *>(P.dst + (goff + (unsigned) (c * nx))) =
                        make_float4(o[c][0], o[c][1], o[c][2], o[c][3]);
                if (want_max) {
#pragma unroll
                    for (int c = 0; c < R; c++) {
                        const float m01 = fmaxf(fabsf(__fsub_rn(pv[c].x, o[c][0])), fabsf(__fsub_rn(pv[c].y, o[c][1])));
                        const float m23 = fmaxf(fabsf(__fsub_rn(pv[c].z, o[c][2])), fabsf(__fsub_rn(pv[c].w, o[c][3])));
                        vmax = fmaxf(vmax, fmaxf(m01, m23));
                    }
                }
                if (dmask != 0 && !(zo & 1) && (zo >> 1) < dnz) {
                    float *qd = P.ds_dst + ((size_t) (zo >> 1) * dny * dnx + (size_t) (x >> 1));
#pragma unroll
                    for (int c = 0; c < R; c++)
                        if (dmask & (1u << c))
                            *reinterpret_cast<float2 *>(qd + (size_t) ((ybase + c) >> 1) * dnx) =
                                make_float2(o[c][0], o[c][2]);
                }
                goff -= plane_stride;
            }
            if (xfix && n + 2 < N) fix(n + 2);
            __syncthreads();
            if (tid == 0 && n + 1 + kNB < N) issue(n + 1 + kNB);
        }
        ebase += (uint32_t) N;
    }  // chunks

    // ---- DoG max-abs: block reduce, one atomic per CTA (non-negative floats order as ints)
    if (P.dogmax != nullptr) {
        for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
        if ((tid & 31) == 0) wmax_sh[tid >> 5] = vmax;
        __syncthreads();
        if (tid < 32) {
            float m2 = tid < NT / 32 ? wmax_sh[tid] : 0.0f;
            for (int o = 16; o > 0; o >>= 1) m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
            if (tid == 0) atomicMax(reinterpret_cast<int *>(P.dogmax), __float_as_int(m2));
        }
    }
}

}  // namespace s3d
