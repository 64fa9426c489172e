// blur_fused.cuh -- the roofline kernel: one launch produces one Gaussian level from the
// previous one (x, y and z passes fused; 1 HBM read + 1 HBM write per voxel), for levels
// whose tap spacing is exactly one voxel on every axis (octave 0 with units (1,1,1)).
//
//   * a CTA owns a TX x TY column of the volume and marches along z through a z-chunk;
//   * every input plane tile (with x/y halo) is staged in shared memory by TMA
//     (cp.async.bulk.tensor.3d + mbarrier, double buffered, issued by one thread);
//   * x pass: shared -> registers (LDS.128) -> shared; y pass: shared -> registers;
//   * z pass: a (2*HW+1)-deep window of y-filtered values lives in registers; the window
//     rotation is resolved at compile time by switching on the phase;
//   * epilogue: max|G(s-1) - G(s)| (the DoG level's max-abs, sift.c:822-826) and the 2x
//     decimation into the next octave (imutil.c:591-617) are fused into the store.
//
// Bit parity with convolve_sep_gen (imutil.c:742-861): taps are accumulated in ascending
// d with separate fp32 multiply and add roundings. Interior samples have frac == 0, so
// (1-frac)*lo + frac*hi is lo itself. The left boundary of a dyadic unit-spaced line is
// an exact reflection (c = -c with integral c), which the z stream reproduces by
// re-feeding planes HW..1; the right boundary's "- 0.1" lerp (imutil.c:846-849) becomes
// HW+1 "virtual planes" built from the last real planes. x and y boundaries evaluate the
// literal per-tap formula (from global memory / from the shared x-filtered rows).
#pragma once

#include <cuda.h>

#include "common.cuh"

namespace s3d {

constexpr int kFTX = 32;  // tile width (x): 8 float4 column groups

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
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *tmap, int x, int y, int z,
                                            uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst)),
        "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
        : "memory");
}

struct FusedParams {
    const float *src;      // previous level (x/y boundary taps, DoG max)
    float *dst;            // this level
    float *ds_dst;         // next octave's level -1 (2x decimation of dst) or nullptr
    float *dogmax;         // slot for max|src - dst| or nullptr
    const float *divisor;  // im_scale divisor applied to src on load, or nullptr
    const float *taps_g;   // taps in global memory (boundary evaluation)
    int nx, ny, nz;
    int zchunk;            // outputs per z-chunk
    TapsParam taps;        // taps in the parameter bank (interior evaluation)
};

template <int HW>
struct FusedCfg {
    static constexpr int NW = 2 * HW + 1;           // z window depth
    static constexpr int HWP = (HW + 3) / 4 * 4;    // x halo rounded up for 16-byte alignment
    static constexpr int BW = kFTX + 2 * HWP;       // staged tile width
    static constexpr int NIN = 4 + 2 * HWP;         // inputs a 4-wide x item reads
};

// One z step with the newest y-filtered value in window slot J (compile-time rotation).
template <int HW, int J>
__device__ __forceinline__ float4 zstep(float4 (&W)[2 * HW + 1], const float4 v, const TapsParam &k) {
    constexpr int NW = 2 * HW + 1;
    W[J] = v;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int t = 0; t < NW; t++) {
        const float4 s = W[(J - t + NW) % NW];  // tap d = t - HW samples plane z + HW - t
        const float kt = k.k[t];
        a.x = __fadd_rn(a.x, __fmul_rn(kt, s.x));
        a.y = __fadd_rn(a.y, __fmul_rn(kt, s.y));
        a.z = __fadd_rn(a.z, __fmul_rn(kt, s.z));
        a.w = __fadd_rn(a.w, __fmul_rn(kt, s.w));
    }
    return a;
}

template <int HW, int J>
struct ZDispatch {
    static __device__ __forceinline__ float4 run(int phase, float4 (&W)[2 * HW + 1], const float4 v,
                                                 const TapsParam &k) {
        if (phase == J) return zstep<HW, J>(W, v, k);
        return ZDispatch<HW, J - 1>::run(phase, W, v, k);
    }
};
template <int HW>
struct ZDispatch<HW, 0> {
    static __device__ __forceinline__ float4 run(int, float4 (&W)[2 * HW + 1], const float4 v,
                                                 const TapsParam &k) {
        return zstep<HW, 0>(W, v, k);
    }
};

__device__ __forceinline__ float4 lerp4(const float4 a, const float4 b, const float fr) {
    const float om = __fsub_rn(1.0f, fr);
    float4 r;
    r.x = __fadd_rn(__fmul_rn(om, a.x), __fmul_rn(fr, b.x));
    r.y = __fadd_rn(__fmul_rn(om, a.y), __fmul_rn(fr, b.y));
    r.z = __fadd_rn(__fmul_rn(om, a.z), __fmul_rn(fr, b.z));
    r.w = __fadd_rn(__fmul_rn(om, a.w), __fmul_rn(fr, b.w));
    return r;
}

// Literal per-tap evaluation on a line of x-filtered rows held in shared memory
// (y boundaries). `rows` points at row y = row0; consecutive y are `pitch` floats apart.
__device__ __forceinline__ float4 ypoint_generic(const float *rows, int pitch, int row0, int y,
                                                 int ny, const TapsParam &k, int hw) {
    const int dim_end = ny - 1;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int d = -hw; d <= hw; d++) {
        const float tap = k.k[d + hw];
        float c = __fsub_rn((float) y, (float) d);  // unit_factor == 1
        if ((int) c < 0) {
            c = -c;
        } else if ((int) c >= dim_end) {
            c = __fsub_rn(__fsub_rn(__fmul_rn(2.0f, (float) dim_end), c), 0.1f);
        }
        int lo = (int) c;
        const float fr = __fsub_rn(c, (float) lo);
        lo = max(0, min(lo, ny - 2));
        const float4 a = *reinterpret_cast<const float4 *>(rows + (lo - row0) * pitch);
        const float4 b = *reinterpret_cast<const float4 *>(rows + (lo + 1 - row0) * pitch);
        const float4 v = lerp4(a, b, fr);
        acc.x = __fadd_rn(acc.x, __fmul_rn(tap, v.x));
        acc.y = __fadd_rn(acc.y, __fmul_rn(tap, v.y));
        acc.z = __fadd_rn(acc.z, __fmul_rn(tap, v.z));
        acc.w = __fadd_rn(acc.w, __fmul_rn(tap, v.w));
    }
    return acc;
}

// Fused blur, unit tap spacing on all three axes. Block = 8 x TY threads; thread
// (g, yy) owns the four columns x0 + 4g .. 4g+3 of row y0 + yy in the y and z passes.
template <int HW, int TY>
__global__ void __launch_bounds__(8 * TY, 1)
blur_fused_unit_kernel(const __grid_constant__ CUtensorMap tmap, const FusedParams P) {
    using C = FusedCfg<HW>;
    constexpr int NT = 8 * TY;
    constexpr int NW = C::NW, HWP = C::HWP, BW = C::BW, NIN = C::NIN;
    constexpr int TROWS = TY + 2 * HW;            // staged rows (y halo)
    constexpr int TILE_FLOATS = TROWS * BW;
    constexpr uint32_t TILE_BYTES = TILE_FLOATS * 4;
    constexpr int TILE_STRIDE = (TILE_FLOATS + 31) / 32 * 32;  // keeps both TMA destinations 128-byte aligned

    extern __shared__ __align__(128) unsigned char smem_raw[];
    float *tile = reinterpret_cast<float *>(smem_raw);                 // [2][TROWS][BW]
    float *xf = tile + 2 * TILE_STRIDE;                                // [2][TROWS][kFTX]
    uint64_t *mbar = reinterpret_cast<uint64_t *>(xf + 2 * TROWS * kFTX);  // [2]
    __shared__ float wmax_sh[NT / 32];

    const int tid = threadIdx.x;
    const int g = tid & 7, yy = tid >> 3;
    const int nx = P.nx, ny = P.ny, nz = P.nz;
    // tiles are shifted inwards at the high edges so every tile is full (overlapping
    // tiles write identical values)
    const int x0 = min((int) blockIdx.x * kFTX, nx - kFTX);
    const int y0 = min((int) blockIdx.y * TY, ny - TY);
    const int zb = blockIdx.z * P.zchunk;
    const int ze = min(zb + P.zchunk, nz);
    const bool last = ze == nz;
    const int dim_end = nz - 1;
    // stream of plane indices i = zb-HW .. i_last; negative i reflect (plane -i)
    const int i0 = zb - HW;
    const int i_last = last ? dim_end - 1 : ze - 1 + HW;
    const int nmain = i_last - i0 + 1;
    const int nloads = nmain + (last ? 1 : 0);  // the last chunk also stages plane nz-1

    float dv = 0.0f;
    if (P.divisor != nullptr) dv = *P.divisor;
    const bool scale = dv != 0.0f;

    if (tid == 0) {
        mbar_init(&mbar[0], 1);
        mbar_init(&mbar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int n) {  // stage stream element n into buffer n & 1 (one thread)
        const int i = i0 + n;
        const int plane = n < nmain ? (i < 0 ? -i : i) : dim_end;
        mbar_expect_tx(&mbar[n & 1], TILE_BYTES);
        tma_load_3d(tile + (n & 1) * TILE_STRIDE, &tmap, x0 - HWP, y0 - HW, plane, &mbar[n & 1]);
    };
    if (tid == 0) {
        issue(0);
        if (nloads > 1) issue(1);
    }

    float4 W[NW];
#pragma unroll
    for (int j = 0; j < NW; j++) W[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    int phase = last ? (NW - nmain % NW) % NW : 0;  // so that plane nz-2 lands in slot NW-1
    float vmax = 0.0f;
    const int y = y0 + yy;
    const bool y_interior = y >= HW && y <= ny - 2 - HW;
    const size_t plane_stride = (size_t) nx * ny;

    // Stores one output row segment (+ fused DoG max and 2x decimation).
    auto emit = [&](int z, const float4 o) {
        const size_t idx = (size_t) z * plane_stride + (size_t) y * nx + (size_t) (x0 + 4 * g);
        *reinterpret_cast<float4 *>(P.dst + idx) = o;
        if (P.dogmax != nullptr) {
            const float4 p = *reinterpret_cast<const float4 *>(P.src + idx);
            vmax = fmaxf(vmax, fmaxf(fmaxf(fabsf(__fsub_rn(p.x, o.x)), fabsf(__fsub_rn(p.y, o.y))),
                                     fmaxf(fabsf(__fsub_rn(p.z, o.z)), fabsf(__fsub_rn(p.w, o.w)))));
        }
        if (P.ds_dst != nullptr && !((z | y) & 1)) {
            const int dnx = nx >> 1, dny = ny >> 1, dnz = nz >> 1;
            const int hz = z >> 1, hy = y >> 1, hx = (x0 + 4 * g) >> 1;
            if (hz < dnz && hy < dny) {
                float *q = P.ds_dst + ((size_t) hz * dny + hy) * dnx + hx;
                if (hx < dnx) q[0] = o.x;
                if (hx + 1 < dnx) q[1] = o.z;
            }
        }
    };

    // x pass + y pass of stream element n; returns this thread's y-filtered float4.
    auto xy_pass = [&](int n, int plane) -> float4 {
        float *tb = tile + (n & 1) * TILE_STRIDE;
        float *xb = xf + (n & 1) * (TROWS * kFTX);
        mbar_wait(&mbar[n & 1], (uint32_t) ((n >> 1) & 1));
        if (scale) {  // im_scale: v / max on load (imutil.c:710-712), first blur only
            for (int e = tid; e < TILE_FLOATS; e += NT) tb[e] = __fdiv_rn(tb[e], dv);
            __syncthreads();
        }
        // ---- x pass: items = (row, 4-column group)
        for (int it = tid; it < TROWS * 8; it += NT) {
            const int r = it >> 3, gg = it & 7;
            const int gy = y0 - HW + r;
            if (gy < 0 || gy >= ny) continue;
            const int gx = x0 + 4 * gg;
            float o[4];
            if (gx >= HW && gx + 3 <= nx - 2 - HW) {
                float in[NIN];
                const float4 *src4 = reinterpret_cast<const float4 *>(tb + r * BW + 4 * gg);
#pragma unroll
                for (int q = 0; q < NIN / 4; q++) {
                    const float4 t4 = src4[q];
                    in[4 * q] = t4.x; in[4 * q + 1] = t4.y; in[4 * q + 2] = t4.z; in[4 * q + 3] = t4.w;
                }
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    float a = 0.0f;
#pragma unroll
                    for (int t = 0; t < NW; t++)  // d = t - HW samples x - d
                        a = __fadd_rn(a, __fmul_rn(P.taps.k[t], in[c + HWP + HW - t]));
                    o[c] = a;
                }
            } else {
                const float *line = P.src + (size_t) plane * plane_stride + (size_t) gy * nx;
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const int x = gx + c;
                    if (x >= HW && x <= nx - 2 - HW) {
                        float a = 0.0f;
                        for (int t = 0; t < NW; t++) {
                            float s = line[x + HW - t];
                            if (scale) s = __fdiv_rn(s, dv);
                            a = __fadd_rn(a, __fmul_rn(P.taps.k[t], s));
                        }
                        o[c] = a;
                    } else {
                        o[c] = conv_point_generic(line, 1, x, nx, P.taps_g, HW, 1.0f, HW, scale, dv);
                    }
                }
            }
            *reinterpret_cast<float4 *>(xb + r * kFTX + 4 * gg) = make_float4(o[0], o[1], o[2], o[3]);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        // the staged tile is consumed: refill this buffer with stream element n + 2
        if (tid == 0 && n + 2 < nloads) issue(n + 2);
        // ---- y pass: thread (g, yy), rows yy .. yy + 2*HW of the x-filtered tile
        float4 v;
        if (y_interior) {
            v = make_float4(0.f, 0.f, 0.f, 0.f);
            const float *col = xb + yy * kFTX + 4 * g;
#pragma unroll
            for (int t = 0; t < NW; t++) {  // d = t - HW samples row y - d = yy + 2*HW - t
                const float4 s = *reinterpret_cast<const float4 *>(col + (2 * HW - t) * kFTX);
                const float kt = P.taps.k[t];
                v.x = __fadd_rn(v.x, __fmul_rn(kt, s.x));
                v.y = __fadd_rn(v.y, __fmul_rn(kt, s.y));
                v.z = __fadd_rn(v.z, __fmul_rn(kt, s.z));
                v.w = __fadd_rn(v.w, __fmul_rn(kt, s.w));
            }
        } else {
            v = ypoint_generic(xb + 4 * g, kFTX, y0 - HW, y, ny, P.taps, HW);
        }
        return v;
    };

    // ---- main stream
    for (int n = 0; n < nmain; n++) {
        const int i = i0 + n;
        const float4 v = xy_pass(n, i < 0 ? -i : i);
        const float4 o = ZDispatch<HW, NW - 1>::run(phase, W, v, P.taps);
        phase = phase + 1 == NW ? 0 : phase + 1;
        const int z = i - HW;
        if (z >= zb && z < ze) emit(z, o);
    }

    // ---- right boundary (last chunk): plane nz-1 and HW "virtual planes"
    if (last) {
        const float4 vreal = xy_pass(nmain, dim_end);
#pragma unroll
        for (int m = 0; m <= HW; m++) {
            // tap position c = dim_end + m mirrors to 2*dim_end - c - 0.1 (imutil.c:846-849)
            const float c = (float) (dim_end + m);
            const float cm = __fsub_rn(__fsub_rn(__fmul_rn(2.0f, (float) dim_end), c), 0.1f);
            const float fr = __fsub_rn(cm, (float) (int) cm);  // (int) cm == dim_end - m - 1
            // plane dim_end-1-q sits in slot NW-1-q; slots 0..m-1 already hold virtual planes
            const float4 a = W[NW - 1 - m];
            const float4 b = m == 0 ? vreal : W[NW - m];
            const float4 vm = lerp4(a, b, fr);
            // slot m receives virtual plane m; the window rotation is static here
            const float4 o = ZDispatch<HW, NW - 1>::run(m, W, vm, P.taps);
            emit(dim_end + m - HW, o);
        }
    }

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
