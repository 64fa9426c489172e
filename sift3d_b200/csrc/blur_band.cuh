// blur_band.cuh -- fast-mode blur for ANY tap spacing (anisotropic units, octaves >= 2): every
// pass of convolve_sep_gen (imutil.c:742-861) is a linear map of a line onto itself,
//     out[p] = sum_d k[d] * ((1 - fr) * in[lo] + fr * in[lo + 1]),   (lo, fr) = the literal
//     coordinate rule for (p, d): carried interior chain, left mirror, "- 0.1" right mirror,
// i.e. a BANDED matrix W with out[p] = sum_j W[p][j] * in[base[p] + j] and a band of
// 2 * uhw + 3 samples. The host walks the literal rule once per line position (same float
// operations as conv_point_generic, so every weight lands on the sample the reference reads)
// and folds the 2 * (2 hw + 1) products of a position into its band; the kernels are then
// plain multiply-adds over the band: ~12 - 32 FFMA per voxel and axis instead of 17 taps x
// (two loads, a float -> int -> float round trip, five unfused operations), at a reordering of
// the sum (~1e-7 of the level's max-abs; the contract is 1e-5). The exact mode keeps the literal
// kernels of blur.cu.
#pragma once

#include <string.h>

#include <mutex>
#include <vector>

#include "common.cuh"

namespace s3d {

// Device form of an axis: positions in blocks of R (4 along x, 8 along y and z), block k holding
// positions R k .. R k + R - 1,
//     out[R k + i] = sum_r w[(r * nblk + k) * R + i] * in[bb[k] + r],   r < BB,
// the union of the R bands (zero weights where a position does not reach): a thread computes
// the R positions of a block from ONE pass over the BB samples -- 32 multiply-adds per 16-byte
// load for the y and z passes -- instead of re-reading the band per position.
struct BandAxis {
    const float4 *w4;  // [BB][nblk][R / 4]
    const int *bb;     // [nblk]: first sample of block k (array coordinates; a multiple of 4 for the x axis)
    int BB, nblk, n;
};

// ---- host: the band of one axis -----------------------------------------------------------
struct BandKey {
    int device, n_line, org, hw;
    float uf;
    float taps[SIFT3D_CUDA_MAX_TAPS];
};
struct BandEntry {
    BandKey key;
    float4 *d_w = nullptr;
    int *d_bb = nullptr;
    int BB = 0, nblk = 0, n = 0;
    int align = 1, R = 4;
};

// Literal coordinate rule of one (position, tap): array sample lo (clamped like samp()) and fraction.
static inline void band_rule(int p, int d, float uf, int n_line, int org, int uhw, float &carried, int &lo,
                             float &fr) {
    const int start = uhw, end = n_line - 2 - uhw, dim_end = n_line - 1;
    float c;
    if (p >= start && p <= end) {  // interior pass: the coordinate is carried across taps (imutil.c:803-818)
        const float step = (float) d * uf;
        carried = carried - step;
        c = carried;
        carried = carried + step;
    } else {  // boundary pass (imutil.c:833-854)
        const float step = (float) d * uf;
        c = (float) p - step;
        if ((int) c < 0) c = -c;
        else if ((int) c >= dim_end) c = 2.0f * (float) dim_end - c - 0.1f;
    }
    lo = (int) c;
    fr = c - (float) lo;
    lo = lo > n_line - 2 ? n_line - 2 : lo;
    lo = (lo < org ? org : lo) - org;
}

static int build_band(const BandKey &K, std::vector<float> &wT, std::vector<int> &base, int &B,
                      std::vector<int> *lo_min_out = nullptr, std::vector<int> *lo_max_out = nullptr) {
    const int n = K.n_line - K.org;  // samples held by the array
    const int uhw = (int) ceilf((float) K.hw * K.uf);
    if (n < 2) return -1;
    std::vector<std::vector<std::pair<int, float>>> rows((size_t) n);
    int width = 1;
    std::vector<int> lo_min((size_t) n), lo_max((size_t) n);
    for (int pa = 0; pa < n; pa++) {
        const int p = pa + K.org;
        float carried = (float) p;
        int mn = n, mx = -1;
        auto &row = rows[(size_t) pa];
        for (int d = -K.hw; d <= K.hw; d++) {
            int lo;
            float fr;
            band_rule(p, d, K.uf, K.n_line, K.org, uhw, carried, lo, fr);
            const float tap = K.taps[d + K.hw];
            const float w0 = tap * (1.0f - fr), w1 = tap * fr;
            row.push_back(std::make_pair(lo, w0));
            mn = lo < mn ? lo : mn;
            mx = lo > mx ? lo : mx;
            if (fr != 0.0f) {  // (the reference reads in[lo + 1] times 0 otherwise)
                row.push_back(std::make_pair(lo + 1, w1));
                mx = lo + 1 > mx ? lo + 1 : mx;
            }
        }
        lo_min[(size_t) pa] = mn;
        lo_max[(size_t) pa] = mx;
        width = mx - mn + 1 > width ? mx - mn + 1 : width;
    }
    B = width < n ? width : n;
    base.assign((size_t) n, 0);
    wT.assign((size_t) B * n, 0.0f);
    for (int pa = 0; pa < n; pa++) {
        int b = lo_min[(size_t) pa];
        if (b + B > n) b = n - B;
        if (b < 0) b = 0;
        if (lo_max[(size_t) pa] - b >= B) return -1;
        base[(size_t) pa] = b;
        for (const auto &e : rows[(size_t) pa]) wT[(size_t) (e.first - b) * n + pa] += e.second;
    }
    if (lo_min_out) *lo_min_out = lo_min;
    if (lo_max_out) *lo_max_out = lo_max;
    return 0;
}

// Blocks of R positions (the device form above). `align`: bb[] and BB are multiples of it.
static int pack_band_blocks(int n, int B, const std::vector<float> &wT, const std::vector<int> &base,
                            const std::vector<int> &lo_min, const std::vector<int> &lo_max, int align, int R,
                            std::vector<float> &w4, std::vector<int> &bb, int &BB) {
    const int nblk = (n + R - 1) / R;
    bb.assign((size_t) nblk, 0);
    int need = 1;
    for (int k = 0; k < nblk; k++) {
        int mn = n, mx = -1;
        for (int p = R * k; p < R * k + R && p < n; p++) {
            mn = lo_min[(size_t) p] < mn ? lo_min[(size_t) p] : mn;
            mx = lo_max[(size_t) p] > mx ? lo_max[(size_t) p] : mx;
        }
        mn = mn / align * align;
        bb[(size_t) k] = mn;
        need = mx - mn + 1 > need ? mx - mn + 1 : need;
    }
    BB = (need + align - 1) / align * align;
    if (BB > n) {
        if (n % align) return -1;
        BB = n;
    }
    for (int k = 0; k < nblk; k++) {
        if (bb[(size_t) k] + BB > n) bb[(size_t) k] = (n - BB) / align * align;
        if (bb[(size_t) k] < 0 || bb[(size_t) k] + BB > n) return -1;
    }
    w4.assign((size_t) BB * nblk * R, 0.0f);
    for (int p = 0; p < n; p++) {
        const int k = p / R, i = p % R;
        for (int j = 0; j < B; j++) {
            const float w = wT[(size_t) j * n + p];
            if (w == 0.0f) continue;
            const int r = base[(size_t) p] + j - bb[(size_t) k];
            if (r < 0 || r >= BB) return -1;
            w4[((size_t) r * nblk + k) * R + i] = w;
        }
    }
    return 0;
}

// Device copy of an axis band, built at first use and cached for the life of the process (a few
// tens of KB per (axis length, spacing, filter)); the upload is synchronous, once.
static int get_band(int n_line, int org, float uf, const sift3d_cuda_filter &f, int align, int R, BandAxis &out) {
    static std::mutex mu;
    static std::vector<BandEntry> cache;
    BandKey K;
    memset(&K, 0, sizeof(K));
    if (cudaGetDevice(&K.device) != cudaSuccess) return -1;
    K.n_line = n_line;
    K.org = org;
    K.hw = f.width / 2;
    K.uf = uf;
    for (int i = 0; i < f.width; i++) K.taps[i] = f.taps[i];
    std::lock_guard<std::mutex> lock(mu);
    for (const BandEntry &e : cache)
        if (e.align == align && e.R == R && memcmp(&e.key, &K, sizeof(K)) == 0) {
            out.w4 = e.d_w; out.bb = e.d_bb; out.BB = e.BB; out.nblk = e.nblk; out.n = e.n;
            return 0;
        }
    if (cache.size() >= 4096) return 1;  // cache full (thousands of shapes in one process): the literal kernels take it
    std::vector<float> wT, w4;
    std::vector<int> base, bb, lo_min, lo_max;
    BandEntry e;
    e.key = K;
    e.align = align;
    e.R = R;
    e.n = n_line - org;
    e.nblk = (e.n + R - 1) / R;
    int B = 0;
    // not representable: the literal kernels take it
    if (build_band(K, wT, base, B, &lo_min, &lo_max)) return 1;
    if (pack_band_blocks(e.n, B, wT, base, lo_min, lo_max, align, R, w4, bb, e.BB)) return 1;
    S3D_CK(cudaMalloc((void **) &e.d_w, w4.size() * sizeof(float)));
    S3D_CK(cudaMalloc((void **) &e.d_bb, bb.size() * sizeof(int)));
    S3D_CK(cudaMemcpy(e.d_w, w4.data(), w4.size() * sizeof(float), cudaMemcpyHostToDevice));
    S3D_CK(cudaMemcpy(e.d_bb, bb.data(), bb.size() * sizeof(int), cudaMemcpyHostToDevice));
    cache.push_back(e);
    out.w4 = e.d_w; out.bb = e.d_bb; out.BB = e.BB; out.nblk = e.nblk; out.n = e.n;
    return 0;
}

// ---- kernels ---------------------------------------------------------------------------------
// x pass: thread = one block of four x positions on ROWS rows. Lanes are adjacent blocks: the
// 16-byte sample loads (bb[] is a multiple of 4) and the 16-byte weight loads are coalesced; the
// weights of a block are shared by its ROWS rows. im_scale (first blur) is folded into the store.
template <int ROWS>
__global__ void __launch_bounds__(128)
band_x_kernel(const float *__restrict__ src, float *__restrict__ dst, int nx, int ny, const BandAxis A,
              const float *__restrict__ divisor, int zoff) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = ROWS * (blockIdx.y * blockDim.y + threadIdx.y);
    const int z = blockIdx.z + zoff;
    if (k >= A.nblk || y >= ny) return;
    const size_t row = ((size_t) z * ny + (size_t) y) * nx;
    const float *__restrict__ q = src + row + A.bb[k];
    const float4 *__restrict__ w = A.w4 + k;
    float4 acc[ROWS];
#pragma unroll
    for (int i = 0; i < ROWS; i++) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < A.BB; r += 4) {
        const float4 w0 = __ldg(w + (size_t) r * A.nblk);
        const float4 w1 = __ldg(w + (size_t) (r + 1) * A.nblk);
        const float4 w2 = __ldg(w + (size_t) (r + 2) * A.nblk);
        const float4 w3 = __ldg(w + (size_t) (r + 3) * A.nblk);
#pragma unroll
        for (int i = 0; i < ROWS; i++) {
            const float4 v = *reinterpret_cast<const float4 *>(q + (size_t) i * nx + r);
            acc[i].x = fmaf(w3.x, v.w, fmaf(w2.x, v.z, fmaf(w1.x, v.y, fmaf(w0.x, v.x, acc[i].x))));
            acc[i].y = fmaf(w3.y, v.w, fmaf(w2.y, v.z, fmaf(w1.y, v.y, fmaf(w0.y, v.x, acc[i].y))));
            acc[i].z = fmaf(w3.z, v.w, fmaf(w2.z, v.z, fmaf(w1.z, v.y, fmaf(w0.z, v.x, acc[i].z))));
            acc[i].w = fmaf(w3.w, v.w, fmaf(w2.w, v.z, fmaf(w1.w, v.y, fmaf(w0.w, v.x, acc[i].w))));
        }
    }
    float s = 1.0f;
    if (divisor != nullptr) {
        const float dv = *divisor;
        if (dv != 0.0f) s = __fdiv_rn(1.0f, dv);  // an all-zero image is left alone (imutil.c:706)
    }
    float *o = dst + row + 4 * (size_t) k;
#pragma unroll
    for (int i = 0; i < ROWS; i++) {
        if (divisor != nullptr) { acc[i].x *= s; acc[i].y *= s; acc[i].z *= s; acc[i].w *= s; }
        *reinterpret_cast<float4 *>(o + (size_t) i * nx) = acc[i];
    }
}

// y or z pass: thread = four x-adjacent voxels (16-byte accesses) on the R = 8 positions of one
// block of the axis; the weights are uniform across the warp. The z pass (EPI) carries the
// epilogue of a level: max|prev - out| and the 2x decimation. Positions [p0, p1) are stored.
constexpr int kBandDepth = 8;      // cp.async ring slots per thread (y and z passes)
constexpr int kBandThreads = 256;  // threads of a y / z pass CTA (the ring's slot pitch)
constexpr int kBandR = 4;          // positions per thread along y / z (8 measured slower twice: fewer, longer threads)
template <int AXIS, bool EPI>
__global__ void __launch_bounds__(256)
band_blk_kernel(const float *__restrict__ src, float *__restrict__ dst, int nx, int ny, int nz, const BandAxis A,
                const float *__restrict__ dog_prev, float *__restrict__ dogmax, float *__restrict__ ds_dst,
                int zoff, int p0, int p1) {
    static_assert(AXIS == 1 || AXIS == 2, "the x pass has a kernel of its own");
    extern __shared__ __align__(16) float4 band_ring[];  // [kBandDepth][threads of the CTA]
    constexpr int R = kBandR;
    const int x = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int u = blockIdx.y * blockDim.y + threadIdx.y;  // AXIS 1: block of y; AXIS 2: y
    const int v_ = blockIdx.z + zoff;                     // AXIS 1: z; AXIS 2: block of z
    const int k = AXIS == 1 ? u : v_;
    float vmax = 0.0f;
    if (x < nx && k < A.nblk && (AXIS == 1 || u < ny)) {
        const size_t stride = AXIS == 1 ? (size_t) nx : (size_t) nx * ny;
        const size_t line = AXIS == 1 ? (size_t) v_ * ny * nx + (size_t) x : (size_t) u * nx + (size_t) x;
        const float *__restrict__ q = src + line + (size_t) A.bb[k] * stride;
        const float4 *__restrict__ w = A.w4 + (size_t) k * (R / 4);
        const size_t wstep = (size_t) A.nblk * (R / 4);
        float4 a[R];
#pragma unroll
        for (int i = 0; i < R; i++) a[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        // The passes are latency-bound (every sample of a thread is a different row or plane): the
        // samples travel global -> shared memory by cp.async into a ring of kBandDepth 16-byte
        // slots PRIVATE to the thread (no barriers; lanes are 16 bytes apart: conflict-free), so
        // kBandDepth - 1 loads per thread are in flight without holding registers. The loop is
        // unrolled over the ring (static slot offsets, advancing pointers): with the loads out of
        // the way the passes are bound by instruction issue.
        constexpr int D = kBandDepth;
        const int BB = A.BB;
        float4 *ring = band_ring + (threadIdx.y * blockDim.x + threadIdx.x);
        const unsigned ring_s = (unsigned) __cvta_generic_to_shared(ring);
        const float *qi = q;  // next sample to request
        int ri = 0;
        auto issue = [&](int slot) {  // an empty group past the band keeps the group count in step
            if (ri < BB)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_s + (unsigned) (slot * kBandThreads * 16)),
                             "l"(qi) : "memory");
            asm volatile("cp.async.commit_group;" ::: "memory");
            qi += stride;
            ri++;
        };
#pragma unroll
        for (int u = 0; u < D - 1; u++) issue(u);
        float4 wcur[R / 4], wnext[R / 4];
#pragma unroll
        for (int h = 0; h < R / 4; h++) wcur[h] = __ldg(w + h);
        const float4 *wp = w;
        for (int r0 = 0; r0 < BB; r0 += D) {
#pragma unroll
            for (int u = 0; u < D; u++) {
                // (the weights of the next round are requested before the wait: the asm statements
                // are barriers for the compiler's own scheduling of loads)
                if (r0 + u + 1 < BB) wp += wstep;
#pragma unroll
                for (int h = 0; h < R / 4; h++) wnext[h] = __ldg(wp + h);
                issue((u + D - 1) % D);  // into the slot read one round ago
                asm volatile("cp.async.wait_group %0;" ::"n"(D - 1) : "memory");
                if (r0 + u < BB) {  // (uniform)
                    const float4 v = ring[u * kBandThreads];
#pragma unroll
                    for (int h = 0; h < R / 4; h++) {
                        const float ws[4] = {wcur[h].x, wcur[h].y, wcur[h].z, wcur[h].w};
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            float4 &t = a[4 * h + i];
                            t.x = fmaf(ws[i], v.x, t.x); t.y = fmaf(ws[i], v.y, t.y);
                            t.z = fmaf(ws[i], v.z, t.z); t.w = fmaf(ws[i], v.w, t.w);
                        }
                    }
                }
#pragma unroll
                for (int h = 0; h < R / 4; h++) wcur[h] = wnext[h];
            }
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int p = R * k + i;
            if (p < p0 || p >= p1) continue;
            const size_t idx = line + (size_t) p * stride;
            const float4 o4 = a[i];
            *reinterpret_cast<float4 *>(dst + idx) = o4;
            if (EPI) {
                const int y = u, z = p;
                if (dogmax != nullptr) {
                    const float4 pq = *reinterpret_cast<const float4 *>(dog_prev + idx);
                    vmax = fmaxf(vmax, fmaxf(fmaxf(fabsf(__fsub_rn(pq.x, o4.x)), fabsf(__fsub_rn(pq.y, o4.y))),
                                             fmaxf(fabsf(__fsub_rn(pq.z, o4.z)), fabsf(__fsub_rn(pq.w, o4.w)))));
                }
                if (ds_dst != nullptr && !((y | z) & 1)) {
                    const int dnx = nx >> 1, dny = ny >> 1, dnz = nz >> 1;
                    const int qx = x >> 1, qy = y >> 1, qz = z >> 1;
                    if (qy < dny && qz < dnz) {
                        float *o = ds_dst + ((size_t) qz * dny + qy) * dnx + qx;
                        if (qx < dnx) o[0] = o4.x;
                        if (qx + 1 < dnx) o[1] = o4.z;
                    }
                }
            }
        }
    }
    if (EPI) {
        if (dogmax != nullptr) block_atomic_max(vmax, dogmax);
    }
}

}  // namespace s3d
