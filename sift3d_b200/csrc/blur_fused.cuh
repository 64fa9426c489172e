// blur_fused.cuh -- the roofline kernel: one launch produces one Gaussian level from the
// previous one (x, y and z passes fused; 1 HBM read + 1 HBM write per voxel), for levels
// whose tap spacing is exactly one voxel on every axis (octave 0 with units (1,1,1)).
//
// Structure (v3):
//   * a CTA owns a 32 x TY column of the volume and marches DOWN z through a z-chunk;
//   * every input plane tile (with x/y halo) is staged in shared memory by TMA
//     (cp.async.bulk.tensor.3d + mbarrier, 3 buffers, issued by one thread; out-of-volume
//     halo is zero filled by TMA and never used);
//   * x pass: item = (tile row, 4-column group): LDS.128 -> registers -> STS.128 into a
//     double-buffered x-filtered tile; the x pass of plane n+1 and the y/z pass of plane n
//     run between the same pair of barriers (one __syncthreads per plane);
//   * y pass: thread (x = lane, 4 consecutive rows) gathers 4 + 2*HW rows (LDS.32,
//     conflict free) and produces 4 y-filtered values;
//   * z pass in registers, scatter form: because the march is in DESCENDING z, plane p is
//     tap t of output p - HW + t with t growing as the march proceeds, i.e. the partial sums
//     receive their taps in ascending d exactly like convolve_sep_gen accumulates them.
//     Each arriving value is multiplied by the HW + 1 distinct taps once (the Gaussian is
//     symmetric bit for bit, checked on the host) and added into 2*HW + 1 running sums that
//     shift by one slot per plane (the shift is free: A[t] = A[t-1] + p);
//   * epilogue: max|G(s-1) - G(s)| (the DoG level's max-abs, sift.c:822-826) and the 2x
//     decimation into the next octave (imutil.c:591-617) are fused into the store.
//
// Bit parity with convolve_sep_gen (imutil.c:742-861): taps are accumulated in ascending
// d with separate fp32 multiply and add roundings. With unit spacing every sample has
// frac == 0, so (1-frac)*lo + frac*hi is lo itself, and all three boundary rules reduce to
// a plain FIR over an EXTENDED line: E[-i] = E[i] (mirror, c = -c with integral c) and
// E[dim_end + m] = lerp(E[dim_end-m-1], E[dim_end-m], fr_m), m = 0..HW, the "- 0.1" right
// mirror (imutil.c:846-849). Interior outputs never touch index dim_end, so giving index
// dim_end the virtual value is consistent. The extension is built on the fly: in registers
// for x (the edge tiles sit at static positions), by remapping / blending source rows in
// the x pass for y, and by the plane stream for z.
#pragma once

#include <cuda.h>

#include "common.cuh"

namespace s3d {

constexpr int kFTX = 32;  // tile width (x): one warp lane per column in the y/z passes
constexpr int kNB = 3;    // TMA tile buffers

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

struct FusedParams {
    const float *src;      // previous level (DoG max)
    float *dst;            // this level
    float *ds_dst;         // next octave's level -1 (2x decimation of dst) or nullptr
    float *dogmax;         // slot for max|src - dst| or nullptr
    const float *divisor;  // unused by the fused kernels (im_scale runs as its own pass)
    int nx, ny, nz;
    int z0, z1;            // output planes [z0, z1)
    int zorg;              // z-slab: plane of the whole volume that array plane 0 is (else 0)
    int zchunk;            // outputs per z-chunk
    TapsParam taps;        // taps in the parameter bank
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

// tap index with the symmetric twin folded onto the lower half, so that the compiler sees
// k[t] * v and k[2*HW - t] * v as the same product
template <int HW>
__device__ __forceinline__ constexpr int fold(int t) {
    return t <= HW ? t : 2 * HW - t;
}

// x pass of one tile row for one 4-column group: 4 x-filtered outputs.
//   trow : tile row (column 0 <-> x0 - HWP)        gg : column group 0..7
//   left / right : this tile sits at x0 == 0 / x0 == nx - 32 (static edge positions), so
//                  the register indices of the edge fix-ups are static per column group
//   frx[m] : right-mirror fractions
template <int HW>
__device__ __forceinline__ float4 xrow(const float *__restrict__ trow, const int gg, const bool left,
                                       const bool right, const float *__restrict__ frx,
                                       const TapsParam &k) {
    using C = FusedCfg<HW>;
    constexpr int NW = C::NW, HWP = C::HWP, NIN = C::NIN;
    float in[NIN];
    const float4 *s4 = reinterpret_cast<const float4 *>(trow + 4 * gg);
#pragma unroll
    for (int q = 0; q < NIN / 4; q++) {
        const float4 t4 = s4[q];
        in[4 * q] = t4.x; in[4 * q + 1] = t4.y; in[4 * q + 2] = t4.z; in[4 * q + 3] = t4.w;
    }
    if (left) {  // E[-i] = E[i]: input j of group GG sits at x = 4*GG + j - HWP
#pragma unroll
        for (int GG = 0; GG < (HWP + 3) / 4; GG++) {
            if (gg == GG) {
#pragma unroll
                for (int j = 0; j < HWP - 4 * GG; j++) in[j] = in[2 * (HWP - 4 * GG) - j];
            }
        }
    }
    if (right) {  // E[dim_end + m] = lerp(E[dim_end-m-1], E[dim_end-m]); dim_end <-> input JD
#pragma unroll
        for (int GG = (kFTX - 4 - HWP + 3) / 4; GG < 8; GG++) {
            if (gg == GG) {
#pragma unroll
                for (int m = HW; m >= 0; m--) {
                    if (kFTX - 1 + HWP - 4 * GG + m <= NIN - 1)
                        in[kFTX - 1 + HWP - 4 * GG + m] =
                            lerp1(in[kFTX - 1 + HWP - 4 * GG - m - 1], in[kFTX - 1 + HWP - 4 * GG - m], frx[m]);
                }
            }
        }
    }
    float o[4];
#pragma unroll
    for (int c = 0; c < 4; c++) {
        float a = 0.0f;
#pragma unroll
        for (int t = 0; t < NW; t++)  // tap d = t - HW samples x - d
            a = __fadd_rn(a, __fmul_rn(k.k[fold<HW>(t)], in[c + HWP + HW - t]));
        o[c] = a;
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

// z pass, scatter form: feed one y-filtered value; returns the output that is now complete.
template <int HW>
__device__ __forceinline__ float zfeed(float (&A)[2 * HW], const float v, const TapsParam &k) {
    float P[HW + 1];
#pragma unroll
    for (int j = 0; j <= HW; j++) P[j] = __fmul_rn(k.k[j], v);
    const float out = __fadd_rn(A[2 * HW - 1], P[0]);
#pragma unroll
    for (int t = 2 * HW - 1; t >= 1; t--) A[t] = __fadd_rn(A[t - 1], P[fold<HW>(t)]);
    A[0] = __fadd_rn(0.0f, P[0]);
    return out;
}

// Chunk table: CTA i produces the outputs [b[i], b[i+1]) of the flattened (column tile,
// output plane) space, tile-major: unit u is plane z0 + u % nzr of tile u / nzr. The host
// cuts the space so that every CTA has the same cost (outputs + 2*HW halo planes per chunk,
// a chunk being the part of a CTA's range inside one tile), see fused_chunk_table().
constexpr int kMaxFusedCtas = 3 * kNumSMs;
struct ChunkTable {
    int b[kMaxFusedCtas + 1];
};

// Fused blur, unit tap spacing on all three axes. Block = 32 x TY/4 threads: thread
// (lane, w) owns column x0 + lane, rows y0 + 4w .. 4w+3 in the y and z passes.
template <int HW, int TY>
__global__ void __launch_bounds__(8 * TY, TY == 64 ? 1 : (HW <= 4 ? 3 : 2))
blur_fused_unit_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ FusedParams P,
                       const __grid_constant__ ChunkTable tab) {
    using C = FusedCfg<HW>;
    constexpr int NT = 8 * TY;
    constexpr int HWP = C::HWP, BW = C::BW, NINY = C::NINY;
    constexpr int TROWS = TY + 2 * HW;            // staged rows (y halo)
    constexpr int TILE_FLOATS = TROWS * BW;
    constexpr uint32_t TILE_BYTES = TILE_FLOATS * 4;
    constexpr int TILE_STRIDE = (TILE_FLOATS + 31) / 32 * 32;  // 128-byte aligned TMA destinations
    constexpr int XF_FLOATS = TROWS * kFTX;
    constexpr int NITEMS = TROWS * 8;
    constexpr int ROT = (TROWS - NT / 8 % TROWS) % TROWS;  // see the x-pass items
    // the four extra live registers would spill in the two variants at their register cap
    constexpr bool kPrefetchSrc = HW < 8 && !(TY == 32 && HW == 4);

    extern __shared__ __align__(128) unsigned char smem_raw[];
    float *tile = reinterpret_cast<float *>(smem_raw);                 // [kNB][TROWS][BW]
    float *xf = tile + kNB * TILE_STRIDE;                              // [2][TROWS][32]
    uint64_t *mbar = reinterpret_cast<uint64_t *>(xf + 2 * XF_FLOATS); // [kNB]
    float *frx = reinterpret_cast<float *>(mbar + kNB);                // [HW + 1]
    float *fry = frx + (HW + 1);                                       // [HW + 1]
    __shared__ float wmax_sh[NT / 32];

    const int tid = threadIdx.x;
    const int lane = tid & 31, w = tid >> 5;
    const int gg = tid & 7;
    const int nx = P.nx, ny = P.ny, nz = P.nz;
    const int dex = nx - 1, dey = ny - 1, dez = nz - 1;
    const int nzr = P.z1 - P.z0;
    const int ntx = (nx + kFTX - 1) / kFTX;
    const size_t plane_stride = (size_t) nx * ny;
    const int dnx = nx >> 1, dny = ny >> 1, dnz = nz >> 1;

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

    // Tile buffers are filled and drained in the same cyclic order over the whole kernel
    // (every chunk issues and consumes exactly its N stream elements), so the positions
    // simply carry over from chunk to chunk.
    int ib_buf = 0;          // (thread 0) tile buffer the next TMA load goes to
    int xb_buf = 0;          // tile buffer of the element the x pass consumes next
    uint32_t xb_par = 0;     // and its mbarrier phase parity
    float vmax = 0.0f;

    int pos = tab.b[blockIdx.x];
    const int pend = tab.b[blockIdx.x + 1];
    while (pos < pend) {
        const int tl = pos / nzr;
        const int zoff = pos - tl * nzr;
        const int zlen = min(pend - pos, nzr - zoff);
        pos += zlen;
        const int tyi = tl / ntx, txi = tl - tyi * ntx;
        // tiles are shifted inwards at the high edges so every tile is full
        const int x0 = min(txi * kFTX, nx - kFTX);
        const int y0 = min(tyi * TY, ny - TY);
        const bool xleft = x0 == 0, xright = x0 == nx - kFTX;
        const bool ybot = y0 == 0, ytop = y0 == ny - TY;
        const int zb = P.z0 + zoff;  // outputs [zb, ze) of this tile
        const int ze = zb + zlen;
        const bool top = ze == nz;

        // Stream of fed positions q (z_out = q + HW). Top chunk: a prologue walks planes
        // dez-HW-1 .. dez upwards to build the virtual planes dez+HW .. dez (fed in that
        // order), then the main stream feeds real planes dez-1 downwards. Other chunks:
        // planes ze-1+HW downwards. Negative positions reflect (plane -q).
        const int npro = top ? HW + 2 : 0;
        const int qstart = top ? dez - 1 : ze - 1 + HW;
        const int N = npro + (qstart - (zb - HW) + 1);
        auto plane_of = [&](int n) -> int {
            if (n < npro) return dez - HW - 1 + n;
            const int q = qstart - (n - npro);
            return q < 0 ? -q : q;
        };
        auto issue = [&](int n) {  // stage stream element n (thread 0 only)
            mbar_expect_tx(&mbar[ib_buf], TILE_BYTES);
            tma_load_3d(tile + ib_buf * TILE_STRIDE, &tmap, x0 - HWP, y0 - HW, plane_of(n), &mbar[ib_buf]);
            if (++ib_buf == kNB) ib_buf = 0;
        };
        if (tid == 0) {
            for (int n = 0; n < kNB && n < N; n++) issue(n);
        }

        float A[4][2 * HW];
#pragma unroll
        for (int c = 0; c < 4; c++)
#pragma unroll
            for (int t = 0; t < 2 * HW; t++) A[c][t] = 0.0f;

        // ---- chunk-invariant x-pass work of this thread: up to 2 items (tile row, 4-column
        // group). Item k covers tile row ((tid + k*NT) >> 3) + ROT: the rotation hands the top
        // tile's virtual rows (two source rows each) to threads that own a single item.
        // offA = float offset of the source row in the tile, offB >= 0: result is
        // lerp(row offB, row offA, frI) (the "- 0.1" right mirror, imutil.c:846-849).
        constexpr int kItems = (NITEMS + NT - 1) / NT;
        static_assert(kItems <= 2, "tile has at most 2 x-pass items per thread");
        int offA[kItems], offB[kItems], offD[kItems];
        float frI[kItems];
#pragma unroll
        for (int k = 0; k < kItems; k++) {
            const int it = tid + k * NT;
            int r = (it >> 3) + ROT;
            if (r >= TROWS) r -= TROWS;
            const int gy = y0 - HW + r;
            int ra = r, rb = -1;
            float fr = 0.0f;
            if (gy < 0) {
                ra = r - 2 * gy;  // mirror: row -gy
            } else if (gy >= dey) {
                const int m = gy - dey;  // virtual row dey + m; only the top tile's outputs use it
                if (ytop) {
                    ra = r - 2 * m;
                    rb = r - 2 * m - 1;
                    fr = fry[m];
                } else {
                    ra = -1;
                }
            }
            if (it >= NITEMS) ra = -1;
            offA[k] = ra >= 0 ? ra * BW : -1;
            offB[k] = rb >= 0 ? rb * BW : -1;
            offD[k] = r * kFTX + 4 * gg;
            frI[k] = fr;
        }

        // ---- which outputs this thread stores: edge outputs belong to the tiles at the static
        // edge positions (first / last), see xrow()
        const int x = x0 + lane;
        const bool xok = (xleft || x >= HW) && (xright || x + HW < dex);
        const int ybase = y0 + 4 * w;
        unsigned smask = 0;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const int y = ybase + c;
            if (xok && (ybot || y >= HW) && (ytop || y + HW < dey)) smask |= 1u << c;
        }
        // decimation targets: even z, even rows, even x
        unsigned dmask = 0;
        if (P.ds_dst != nullptr && !(x & 1) && (x >> 1) < dnx) {
#pragma unroll
            for (int c = 0; c < 4; c++)
                if ((smask & (1u << c)) && !((ybase + c) & 1) && ((ybase + c) >> 1) < dny) dmask |= 1u << c;
        }
        const size_t ds_col = (size_t) (ybase >> 1) * dnx + (size_t) (x >> 1);
        // element (ze - 1, ybase, x): the first output stored; one plane down per fed plane
        const size_t first_off = (size_t) (ze - 1) * plane_stride + (size_t) ybase * nx + (size_t) x;
        float *dp = P.dst + first_off;
        const float *sp = P.src + first_off;

        // x pass of stream element e into xf[e & 1]
        auto xpass = [&](const int e) {
            const float *tb = tile + xb_buf * TILE_STRIDE;
            float *xb = xf + (e & 1) * XF_FLOATS;
            mbar_wait(&mbar[xb_buf], xb_par);
            if (++xb_buf == kNB) {
                xb_buf = 0;
                xb_par ^= 1u;
            }
#pragma unroll
            for (int k = 0; k < kItems; k++) {
                if (offA[k] < 0) {
                    if (tid + k * NT < NITEMS)
                        *reinterpret_cast<float4 *>(xb + offD[k]) = make_float4(0.f, 0.f, 0.f, 0.f);
                    continue;
                }
                float4 o = xrow<HW>(tb + offA[k], gg, xleft, xright, frx, P.taps);
                if (offB[k] >= 0) {
                    const float4 a = xrow<HW>(tb + offB[k], gg, xleft, xright, frx, P.taps);
                    const float fr = frI[k];
                    o = make_float4(lerp1(a.x, o.x, fr), lerp1(a.y, o.y, fr), lerp1(a.z, o.z, fr),
                                    lerp1(a.w, o.w, fr));
                }
                *reinterpret_cast<float4 *>(xb + offD[k]) = o;
            }
        };
        // y pass of element n: rows 4w .. 4w+3+2*HW of the x-filtered tile, column `lane`
        auto ypass = [&](const int n, float (&yv)[4]) {
            const float *xb = xf + (n & 1) * XF_FLOATS + (4 * w) * kFTX + lane;
            float in[NINY];
#pragma unroll
            for (int j = 0; j < NINY; j++) in[j] = xb[j * kFTX];
#pragma unroll
            for (int c = 0; c < 4; c++) {
                float a = 0.0f;
#pragma unroll
                for (int t = 0; t < 2 * HW + 1; t++)  // tap d = t - HW samples row y - d
                    a = __fadd_rn(a, __fmul_rn(P.taps.k[fold<HW>(t)], in[c + 2 * HW - t]));
                yv[c] = a;
            }
        };
        auto plane_end = [&](const int n) {
            __syncthreads();
            // the tile buffer of element n+1 was consumed by xpass(n+1): refill it
            if (tid == 0 && n + 1 + kNB < N) issue(n + 1 + kNB);
        };

        xpass(0);
        plane_end(-1);
        int n = 0;
        if (top) {
            // ---- prologue (top chunk): y-filtered planes dez-HW-1 .. dez, ascending; from the
            // second one on, the virtual plane dez + m = lerp(plane dez-m-1, plane dez-m) is fed
            float prev[4] = {0.f, 0.f, 0.f, 0.f};
            for (; n < npro; n++) {
                xpass(n + 1);  // the main stream is never empty
                float yv[4];
                ypass(n, yv);
                if (n > 0) {
                    const float fr = right_frac(dez + P.zorg, HW + 1 - n);
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const float v = lerp1(prev[c], yv[c], fr);
                        (void) zfeed<HW>(A[c], v, P.taps);  // outputs beyond the volume
                    }
                }
#pragma unroll
                for (int c = 0; c < 4; c++) prev[c] = yv[c];
                plane_end(n);
            }
        }
        // ---- main stream: the first fed planes only warm the window up (their outputs lie
        // at or above ze); from then on every plane completes the output one plane lower
        int zo = qstart + HW;  // output completed by element n
        for (; n < N; n++, zo--) {
            // source voxels of the DoG max: requested a whole plane step before they are
            // used (an L2 round trip in front of the FMNMX chain cost 15 us per level)
            float pv[4] = {0.f, 0.f, 0.f, 0.f};
            if (kPrefetchSrc && P.dogmax != nullptr && zo < ze) {
#pragma unroll
                for (int c = 0; c < 4; c++)
                    if (smask & (1u << c)) pv[c] = __ldg(sp + c * nx);
            }
            if (n + 1 < N) xpass(n + 1);
            float yv[4];
            ypass(n, yv);
            float o[4];
#pragma unroll
            for (int c = 0; c < 4; c++) o[c] = zfeed<HW>(A[c], yv[c], P.taps);
            if (zo < ze) {
#pragma unroll
                for (int c = 0; c < 4; c++)
                    if (smask & (1u << c)) dp[c * nx] = o[c];
                if (P.dogmax != nullptr) {
#pragma unroll
                    for (int c = 0; c < 4; c++)
                        if (smask & (1u << c))
                            vmax = fmaxf(vmax, fabsf(__fsub_rn(kPrefetchSrc ? pv[c] : sp[c * nx], o[c])));
                }
                if (dmask != 0 && !(zo & 1) && (zo >> 1) < dnz) {
                    float *qd = P.ds_dst + ((size_t) (zo >> 1) * dny * dnx + ds_col);
#pragma unroll
                    for (int c = 0; c < 4; c++)
                        if (dmask & (1u << c))
                            qd[(((ybase + c) >> 1) - (ybase >> 1)) * dnx] = o[c];
                }
                dp -= plane_stride;
                sp -= plane_stride;
            }
            plane_end(n);
        }
    }  // chunks

    // ---- DoG max-abs: block reduce, one atomic per CTA (non-negative floats order as ints)
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
