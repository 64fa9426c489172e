// blur_fused_half.cuh -- fused x/y/z blur for tap spacing 1/2 voxel on every axis
// (octave 1 of an isotropic unit-spaced volume: resize_Pyramid doubles the units,
// imutil.c:1545-1548, while build_gpyr keeps unit = 1, sift.c:675-690).
//
// Same skeleton as blur_fused.cuh (TMA-staged plane tiles, descending z march, scatter-form
// register accumulators, fused DoG max + decimation), plus one idea: with spacing 1/2 the
// taps of output x sample positions x - d/2, i.e. a plain FIR with the ORIGINAL taps over
// the 2x "upsampled" line U,
//     U[2i] = S(i) = src[i]                    U[2i+1] = S(i + 1/2) = lerp(src[i], src[i+1], 1/2)
//     out(x) = sum_t k[t] * U[2x + HW - t]     (ascending t == ascending d, imutil.c:803-818)
// where S(c) is exactly what convolve_sep_gen samples at coordinate c, boundary rules
// included (imutil.c:833-854):
//     c <= -1      (int) c < 0        -> S(-c)
//     c == -1/2    (int) c == 0       -> lerp(src[0], src[1], -1/2)    (no mirror: extrapolation)
//     c >= dim_end                    -> c' = 2*dim_end - c - 0.1f, lerp(src[(int) c'], .., frac)
// Interior outputs never reach those cases, so one definition of U serves both passes of the
// reference. U is built
//   x: in registers, static indices (edge tiles sit at static positions);
//   y: as an explicit shared-memory tile, every row evaluated with the literal rule above
//      (uniform per row), so the y pass is a plain stride-2 FIR for every tile;
//   z: by the plane stream: each loaded plane p yields the pair (U[2p+1], U[2p]).
#pragma once

#include "blur_zmarch.cuh"

namespace s3d {

template <int HW>
struct HalfCfg {
    static constexpr int NW = 2 * HW + 1;
    static constexpr int UHW = (HW + 1) / 2;            // (int) ceilf(HW * 0.5f)
    static constexpr int UHWP = (UHW + 3) / 4 * 4;      // x halo, multiple of 4 (TMA box alignment)
    static constexpr int BW = kFTX + 2 * UHWP;          // staged tile width
    static constexpr int NIN = 4 + 2 * UHWP;            // samples a 4-wide x item reads
    static constexpr int NU = 2 * NIN - 1;              // their upsampled line
    static constexpr int YH = UHW + 1;                  // y halo rows of the staged tile
    // z accumulators: output z = p + j holds taps up to plane p; j in [JMIN, JMAX]
    static constexpr int JMIN = -(HW / 2);
    static constexpr int JMAX = (HW + 1) / 2;
    static constexpr int NA = JMAX - JMIN + 1;          // == HW + 1
};

// The literal sampling rule on an integer "2c" index (c = u / 2), returning lo and frac.
__device__ __forceinline__ void half_rule(int u, int dim_end, int n, int &lo, float &fr) {
    float c = __fmul_rn((float) u, 0.5f);
    if ((int) c < 0) {
        c = -c;
    } else if ((int) c >= dim_end) {
        c = __fsub_rn(__fsub_rn(__fmul_rn(2.0f, (float) dim_end), c), 0.1f);
    }
    lo = (int) c;
    fr = __fsub_rn(c, (float) lo);
    lo = max(0, min(lo, n - 2));
}

// x pass of one tile row for one 4-column group (spacing 1/2).
template <int HW>
__device__ __forceinline__ float4 xrow_half(const float *__restrict__ trow, const int gg,
                                            const bool left, const bool right,
                                            const float *__restrict__ frx, const TapsParam &k) {
    using C = HalfCfg<HW>;
    constexpr int NW = C::NW, UHWP = C::UHWP, NIN = C::NIN, NU = C::NU;
    float in[NIN];
    const float4 *s4 = reinterpret_cast<const float4 *>(trow + 4 * gg);
#pragma unroll
    for (int q = 0; q < NIN / 4; q++) {
        const float4 t4 = s4[q];
        in[4 * q] = t4.x; in[4 * q + 1] = t4.y; in[4 * q + 2] = t4.z; in[4 * q + 3] = t4.w;
    }
    if (left) {  // S(-i) = S(i): sample j of group GG sits at x = 4*GG + j - UHWP
#pragma unroll
        for (int GG = 0; GG < (UHWP + 3) / 4; GG++) {
            if (gg == GG) {
#pragma unroll
                for (int j = 0; j < UHWP - 4 * GG; j++) in[j] = in[2 * (UHWP - 4 * GG) - j];
            }
        }
    }
    float U[NU];  // U[j] = S(x_first + j/2)
#pragma unroll
    for (int j = 0; j < NU; j++) {
        if (j & 1)
            U[j] = lerp1(in[j >> 1], in[(j >> 1) + 1], 0.5f);
        else
            U[j] = in[j >> 1];
    }
    if (left) {  // S(-1/2) is an extrapolation from samples 0 and 1, not a mirror
#pragma unroll
        for (int GG = 0; GG < (UHWP + 3) / 4; GG++) {
            if (gg == GG && UHWP - 4 * GG >= 1)
                U[2 * (UHWP - 4 * GG) - 1] = lerp1(in[UHWP - 4 * GG], in[UHWP - 4 * GG + 1], -0.5f);
        }
    }
    if (right) {  // S(dim_end + m/2), m = 0..HW; dim_end <-> sample JD of this group
#pragma unroll
        for (int GG = (kFTX - 4 - UHWP + 3) / 4; GG < 8; GG++) {
            if (gg == GG) {
#pragma unroll
                for (int m = 0; m <= HW; m++) {
                    if (2 * (kFTX - 1 + UHWP - 4 * GG) + m <= NU - 1)
                        U[2 * (kFTX - 1 + UHWP - 4 * GG) + m] =
                            lerp1(in[kFTX - 1 + UHWP - 4 * GG - (m >> 1) - 1],
                                  in[kFTX - 1 + UHWP - 4 * GG - (m >> 1)], frx[m]);
                }
            }
        }
    }
    float o[4];
#pragma unroll
    for (int c = 0; c < 4; c++) {
        float a = 0.0f;
#pragma unroll
        for (int t = 0; t < NW; t++)  // out(x) = sum k[t] * U[2x + HW - t]
            a = __fadd_rn(a, __fmul_rn(k.k[fold<HW>(t)], U[2 * (c + UHWP) + HW - t]));
        o[c] = a;
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

// z pass, scatter form, one plane step: feed U[2p+1] = h, then U[2p] = v; returns the
// output z = p + JMAX, which is complete after this step.
template <int HW>
__device__ __forceinline__ float zstep_half(float (&A)[HW + 1], const float h, const float v,
                                            const TapsParam &k) {
    using C = HalfCfg<HW>;
    float out = 0.0f;
#pragma unroll
    for (int j = C::JMAX; j >= C::JMIN; j--) {
        const int t1 = 2 * j + HW - 1, t2 = 2 * j + HW;  // taps that planes' pair feeds to z = p + j
        float a = A[j - C::JMIN];
        if (t1 >= 0 && t1 <= 2 * HW) a = __fadd_rn(a, __fmul_rn(k.k[fold<HW>(t1)], h));
        if (t2 >= 0 && t2 <= 2 * HW) a = __fadd_rn(a, __fmul_rn(k.k[fold<HW>(t2)], v));
        if (j == C::JMAX)
            out = a;
        else
            A[j - C::JMIN + 1] = a;
    }
    A[0] = 0.0f;
    return out;
}

// Fused blur, tap spacing 1/2 on all three axes. Block = 32 x TY/4 threads: thread
// (lane, w) owns column x0 + lane, rows y0 + 4w .. 4w+3 in the y and z passes.
// Three pipeline stages run between one pair of barriers per plane:
//   x pass of stream element n+2  ->  U-row build of element n+1  ->  y/z pass of element n.
template <int HW, int TY>
__global__ void __launch_bounds__(8 * TY, TY == 64 ? 1 : 2)
blur_fused_half_kernel(const __grid_constant__ CUtensorMap tmap, const FusedParams P) {
    using C = HalfCfg<HW>;
    constexpr int NT = 8 * TY;
    constexpr int UHW = C::UHW, UHWP = C::UHWP, BW = C::BW, YH = C::YH;
    constexpr int TROWS = TY + 2 * YH;             // staged rows (y halo)
    constexpr int TILE_FLOATS = TROWS * BW;
    constexpr uint32_t TILE_BYTES = TILE_FLOATS * 4;
    constexpr int TILE_STRIDE = (TILE_FLOATS + 31) / 32 * 32;
    constexpr int XF_FLOATS = TROWS * kFTX;
    constexpr int NUR = 2 * TY + 2 * HW - 1;       // U rows: u = 2*y0 - HW .. 2*(y0+TY-1) + HW
    constexpr int U_FLOATS = NUR * kFTX;
    constexpr int NYU = 7 + 2 * HW;                // U rows a 4-high y item reads

    extern __shared__ __align__(128) unsigned char smem_raw[];
    float *tile = reinterpret_cast<float *>(smem_raw);                 // [kNB][TROWS][BW]
    float *xf = tile + kNB * TILE_STRIDE;                              // [2][TROWS][32]
    float *ut = xf + 2 * XF_FLOATS;                                    // [2][NUR][32]
    uint64_t *mbar = reinterpret_cast<uint64_t *>(ut + 2 * U_FLOATS);  // [kNB]
    float *frx = reinterpret_cast<float *>(mbar + kNB);                // [HW + 1]
    __shared__ float wmax_sh[NT / 32];

    const int tid = threadIdx.x;
    const int lane = tid & 31, w = tid >> 5;
    const int nx = P.nx, ny = P.ny, nz = P.nz;
    const int x0 = min((int) blockIdx.x * kFTX, nx - kFTX);
    const int y0 = min((int) blockIdx.y * TY, ny - TY);
    const bool xleft = x0 == 0, xright = x0 == nx - kFTX;
    const int dex = nx - 1, dey = ny - 1, dez = nz - 1;
    const int zb = P.z0 + blockIdx.z * P.zchunk;  // outputs [P.z0, P.z1): the whole volume, or the
    const int ze = min(zb + P.zchunk, P.z1);       // planes this rank owns in z-slab mode
    const bool top = ze == nz;

    // Stream of plane loads, each followed by one plane step p (feeds U[2p+1], U[2p]):
    //   top chunk: planes dez-UHW-1 .. dez ascending build the virtual steps
    //   p = dez+UHW .. dez (the first load only primes `prev`), then dez-1 downwards;
    //   other chunks: planes ze-1+UHW downwards. Steps below 0 reflect (plane -p).
    const int npro = top ? UHW + 2 : 0;
    const int pstart = top ? dez - 1 : ze - 1 + UHW;
    const int pend = zb - UHW;
    const int N = npro + (pstart - pend + 1);
    auto plane_of = [&](int n) -> int {
        if (n < npro) return dez - UHW - 1 + n;
        const int p = pstart - (n - npro);
        return p < 0 ? -p : p;
    };

    if (tid == 0) {
#pragma unroll
        for (int b = 0; b < kNB; b++) mbar_init(&mbar[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid <= HW) {
        int lo;
        float fr;
        half_rule(2 * dex + tid, dex, nx, lo, fr);
        frx[tid] = fr;
    }
    __syncthreads();
    auto issue = [&](int n) {
        const int b = n % kNB;
        mbar_expect_tx(&mbar[b], TILE_BYTES);
        tma_load_3d(tile + b * TILE_STRIDE, &tmap, x0 - UHWP, y0 - YH, plane_of(n), &mbar[b]);
    };
    if (tid == 0) {
        for (int n = 0; n < kNB && n < N; n++) issue(n);
    }

    float A[4][HW + 1];
#pragma unroll
    for (int c = 0; c < 4; c++)
#pragma unroll
        for (int t = 0; t <= HW; t++) A[c][t] = 0.0f;
    float prev[4] = {0.f, 0.f, 0.f, 0.f};
    float vmax = 0.0f;

    const int x = x0 + lane;
    // x edge outputs belong to the tiles at the static edge positions (see xrow_half)
    const bool xok = (xleft || x >= UHW) && (xright || x + UHW < dex);
    const int ybase = y0 + 4 * w;
    const size_t plane_stride = (size_t) nx * ny;
    const size_t col_off = (size_t) ybase * nx + (size_t) x;
    const int dnx = nx >> 1, dny = ny >> 1, dnz = nz >> 1;
    unsigned dmask = 0;  // decimation targets: even z, even rows, even x
    if (P.ds_dst != nullptr && xok && !(x & 1) && (x >> 1) < dnx) {
#pragma unroll
        for (int c = 0; c < 4; c++)
            if (!((ybase + c) & 1) && ((ybase + c) >> 1) < dny) dmask |= 1u << c;
    }
    const size_t ds_col = (size_t) (ybase >> 1) * dnx + (size_t) (x >> 1);

    // ---- plane-invariant work descriptors of this thread
    // x pass: items (tile row, 4-column group), rows inside the volume only
    constexpr int kXItems = (TROWS * 8 + NT - 1) / NT;
    int xoff[kXItems];  // float offset of the row in the tile, -1: nothing to do
    const int gg = tid & 7;
#pragma unroll
    for (int k = 0; k < kXItems; k++) {
        const int it = tid + k * NT;
        const int r = it >> 3;
        const int gy = y0 - YH + r;
        xoff[k] = (it < TROWS * 8 && gy >= 0 && gy <= dey) ? r : -1;
    }
    // U build: items (U row, 4-column group): U[ur] = lerp(xf[lo], xf[lo + 1], fr) by the literal rule
    constexpr int kUItems = (NUR * 8 + NT - 1) / NT;
    int usrc[kUItems];  // float offset of row lo in xf (+ column group), -1: nothing to do
    float ufr[kUItems];
#pragma unroll
    for (int k = 0; k < kUItems; k++) {
        const int it = tid + k * NT;
        const int ur = it >> 3;
        int lo;
        float fr;
        half_rule(2 * y0 - HW + ur, dey, ny, lo, fr);
        const int r = lo - (y0 - YH);  // tile row of sample lo
        usrc[k] = (it < NUR * 8 && r >= 0 && r + 1 < TROWS) ? r * kFTX + 4 * gg : -1;
        ufr[k] = fr;
    }

    int xb_buf = 0;
    uint32_t xb_par = 0;
    for (int n = -2; n < N; n++) {
        // ---- stage 1: x pass of element n + 2 into xf[(n + 2) & 1] (rows inside the volume)
        if (n + 2 < N) {
            const int e = n + 2;
            const float *tb = tile + xb_buf * TILE_STRIDE;
            float *xb = xf + (e & 1) * XF_FLOATS;
            mbar_wait(&mbar[xb_buf], xb_par);
            if (++xb_buf == kNB) {
                xb_buf = 0;
                xb_par ^= 1u;
            }
#pragma unroll
            for (int k = 0; k < kXItems; k++) {
                if (xoff[k] < 0) continue;
                const float4 o = xrow_half<HW>(tb + xoff[k] * BW, gg, xleft, xright, frx, P.taps);
                *reinterpret_cast<float4 *>(xb + xoff[k] * kFTX + 4 * gg) = o;
            }
        }
        // ---- stage 2: U rows of element n + 1: U[ur] = S_y(u / 2), u = 2*y0 - HW + ur
        if (n + 1 >= 0 && n + 1 < N) {
            const int e = n + 1;
            const float *xb = xf + (e & 1) * XF_FLOATS;
            float *ub = ut + (e & 1) * U_FLOATS;
#pragma unroll
            for (int k = 0; k < kUItems; k++) {
                const int it = tid + k * NT;
                if (it >= NUR * 8) continue;
                float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
                if (usrc[k] >= 0) {
                    const float4 a = *reinterpret_cast<const float4 *>(xb + usrc[k]);
                    const float4 c = *reinterpret_cast<const float4 *>(xb + usrc[k] + kFTX);
                    const float fr = ufr[k];
                    o = make_float4(lerp1(a.x, c.x, fr), lerp1(a.y, c.y, fr), lerp1(a.z, c.z, fr),
                                    lerp1(a.w, c.w, fr));
                }
                *reinterpret_cast<float4 *>(ub + (it >> 3) * kFTX + 4 * gg) = o;
            }
        }
        // ---- stage 3: y pass + z step of element n
        if (n >= 0) {
            const float *ub = ut + (n & 1) * U_FLOATS + (8 * w) * kFTX + lane;
            float in[NYU];
#pragma unroll
            for (int j = 0; j < NYU; j++) in[j] = ub[j * kFTX];
            float yv[4];
#pragma unroll
            for (int c = 0; c < 4; c++) {
                float a = 0.0f;
#pragma unroll
                for (int t = 0; t < 2 * HW + 1; t++)  // out(y) = sum k[t] * U[2y + HW - t]
                    a = __fadd_rn(a, __fmul_rn(P.taps.k[fold<HW>(t)], in[2 * c + 2 * HW - t]));
                yv[c] = a;
            }

            int p;
            bool step = true;
            float fr_h = 0.5f, fr_v = 0.0f;  // h = lerp(lo_h, hi_h, fr_h); v likewise (virtual only)
            int mode;                          // 0: normal, 1: virtual, 2: p == -1, 3: p <= -2
            if (n < npro) {
                const int q = dez - UHW - 1 + n;  // loaded plane; pair (q-1, q)
                p = 2 * dez - q;
                mode = 1;
                if (n == 0) step = false;
                int lo;
                // fractions from the coordinates of the whole volume (z-slab: array plane 0 is
                // plane zorg), so that they round as on a single device
                const int zg = P.zorg;
                half_rule(2 * (p + zg) + 1, dez + zg, nz + zg, lo, fr_h);
                half_rule(2 * (p + zg), dez + zg, nz + zg, lo, fr_v);
            } else {
                p = pstart - (n - npro);
                mode = p >= 0 ? 0 : (p == -1 ? 2 : 3);
                if (mode == 2) fr_h = -0.5f;
            }
            if (step) {
                float o[4];
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const float cur = yv[c];
                    float h, v = cur;
                    if (mode == 0) {
                        h = lerp1(cur, prev[c], 0.5f);
                    } else {
                        h = lerp1(prev[c], cur, fr_h);
                        if (mode == 1) v = lerp1(prev[c], cur, fr_v);
                    }
                    o[c] = zstep_half<HW>(A[c], h, v, P.taps);
                }
                const int z = p + C::JMAX;
                if (z >= zb && z < ze && xok) {
                    const size_t base = (size_t) z * plane_stride + col_off;
                    float *dp = P.dst + base;
#pragma unroll
                    for (int c = 0; c < 4; c++) dp[c * nx] = o[c];
                    if (P.dogmax != nullptr) {
                        const float *sp = P.src + base;
#pragma unroll
                        for (int c = 0; c < 4; c++)
                            vmax = fmaxf(vmax, fabsf(__fsub_rn(sp[c * nx], o[c])));
                    }
                    if (dmask != 0 && !(z & 1) && (z >> 1) < dnz) {
                        float *qd = P.ds_dst + ((size_t) (z >> 1) * dny * dnx + ds_col);
#pragma unroll
                        for (int c = 0; c < 4; c++)
                            if (dmask & (1u << c))
                                qd[(((ybase + c) >> 1) - (ybase >> 1)) * dnx] = o[c];
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < 4; c++) prev[c] = yv[c];
        }
        __syncthreads();
        // tile buffer (n+2) % kNB was consumed by the x pass of element n+2: refill it
        if (tid == 0 && n + 2 >= 0 && n + 2 + kNB < N) issue(n + 2 + kNB);
    }

    if (P.dogmax != nullptr) {
        for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
        if (lane == 0) wmax_sh[w] = vmax;
        __syncthreads();
        if (tid < 32) {
            float m2 = tid < NT / 32 ? wmax_sh[tid] : 0.0f;
            for (int o = 16; o > 0; o >>= 1) m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
            if (tid == 0) atomicMax(reinterpret_cast<int *>(P.dogmax), __float_as_int(m2));
        }
    }
}

}  // namespace s3d
