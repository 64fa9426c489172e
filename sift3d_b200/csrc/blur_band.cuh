// blur_band.cuh -- fast-mode blur for ANY tap spacing (anisotropic units, octaves >= 2): every
// pass of convolve_sep_gen (imutil.c:742-861) is a linear map of a line onto itself,
//     out[p] = sum_d k[d] * ((1 - fr) * in[lo] + fr * in[lo + 1]),   (lo, fr) = the literal
//     coordinate rule for (p, d): carried interior chain, left mirror, "- 0.1" right mirror,
// i.e. a BANDED matrix W with out[p] = sum_j W[p][j] * in[base[p] + j] and a band of
// 2 * uhw + 3 samples. The host walks the literal rule once per line position (same float
// operations as conv_point_generic, so every weight lands on the sample the reference reads)
// and folds the 2 * (2 hw + 1) products of a position into its band; the kernels are then
// plain multiply-adds over the band: 15 - 27 FFMA per voxel and axis instead of 17 taps x
// (two loads, a float -> int -> float round trip, five unfused operations), at a reordering of
// the sum (~1e-7 of the level's max-abs; the contract is 1e-5). The exact mode keeps the literal
// kernels of blur.cu.
#pragma once

#include <string.h>

#include <mutex>
#include <vector>

#include "common.cuh"

namespace s3d {

struct BandAxis {
    const float *wT;  // [B][n]: weight j of position p at wT[j * n + p]
    const int *base;  // [n]: first sample of position p's band (array coordinates)
    int B, n;
};

// ---- host: the band of one axis -----------------------------------------------------------
struct BandKey {
    int device, n_line, org, hw;
    float uf;
    float taps[SIFT3D_CUDA_MAX_TAPS];
};
struct BandEntry {
    BandKey key;
    float *d_w = nullptr;
    int *d_base = nullptr;
    int B = 0, n = 0;
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

static int build_band(const BandKey &K, std::vector<float> &wT, std::vector<int> &base, int &B) {
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
    return 0;
}

// Device copy of an axis band, built at first use and cached for the life of the process (a few
// tens of KB per (axis length, spacing, filter)); the upload is synchronous, once.
static int get_band(int n_line, int org, float uf, const sift3d_cuda_filter &f, BandAxis &out) {
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
        if (memcmp(&e.key, &K, sizeof(K)) == 0) {
            out.wT = e.d_w; out.base = e.d_base; out.B = e.B; out.n = e.n;
            return 0;
        }
    if (cache.size() >= 4096) {
        S3D_ERR("sift3d_cuda: band table cache full\n");
        return -1;
    }
    std::vector<float> wT;
    std::vector<int> base;
    BandEntry e;
    e.key = K;
    e.n = n_line - org;
    if (build_band(K, wT, base, e.B)) return 1;  // not representable: the literal kernels take it
    S3D_CK(cudaMalloc((void **) &e.d_w, wT.size() * sizeof(float)));
    S3D_CK(cudaMalloc((void **) &e.d_base, base.size() * sizeof(int)));
    S3D_CK(cudaMemcpy(e.d_w, wT.data(), wT.size() * sizeof(float), cudaMemcpyHostToDevice));
    S3D_CK(cudaMemcpy(e.d_base, base.data(), base.size() * sizeof(int), cudaMemcpyHostToDevice));
    cache.push_back(e);
    out.wT = e.d_w; out.base = e.d_base; out.B = e.B; out.n = e.n;
    return 0;
}

// ---- kernels ---------------------------------------------------------------------------------
// x pass: thread = one column x of four rows (the band of a position depends on x only; lanes are
// adjacent x, so weights and samples are read coalesced). `oscale`: im_scale folded in (first blur).
__global__ void __launch_bounds__(256)
band_x_rows4_kernel(const float *__restrict__ src, float *__restrict__ dst, int nx, int ny, const BandAxis A,
                    const float *__restrict__ divisor, int zoff) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = 4 * (blockIdx.y * blockDim.y + threadIdx.y);
    const int z = blockIdx.z + zoff;
    if (x >= nx || y >= ny) return;
    const size_t row = ((size_t) z * ny + (size_t) y) * nx;
    const float *__restrict__ q = src + row + A.base[x];
    const float *__restrict__ w = A.wT + x;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    for (int j = 0; j < A.B; j++) {
        const float wj = __ldg(w + (size_t) j * nx);
        a0 = fmaf(wj, q[j], a0);
        a1 = fmaf(wj, q[j + (size_t) nx], a1);
        a2 = fmaf(wj, q[j + 2 * (size_t) nx], a2);
        a3 = fmaf(wj, q[j + 3 * (size_t) nx], a3);
    }
    if (divisor != nullptr) {
        const float dv = *divisor;
        if (dv != 0.0f) {  // an all-zero image is left alone (imutil.c:706)
            const float s = __fdiv_rn(1.0f, dv);
            a0 *= s; a1 *= s; a2 *= s; a3 *= s;
        }
    }
    float *o = dst + row + x;
    o[0] = a0;
    o[(size_t) nx] = a1;
    o[2 * (size_t) nx] = a2;
    o[3 * (size_t) nx] = a3;
}

// y or z pass: thread = four x-adjacent voxels (16-byte accesses); the band depends on (y | z) only.
// The z pass (EPI) carries the epilogue of a level: max|prev - out| and the 2x decimation.
template <int AXIS, bool EPI>
__global__ void __launch_bounds__(256)
band4_kernel(const float *__restrict__ src, float *__restrict__ dst, int nx, int ny, int nz, const BandAxis A,
             const float *__restrict__ dog_prev, float *__restrict__ dogmax, float *__restrict__ ds_dst,
             int zoff) {
    static_assert(AXIS == 1 || AXIS == 2, "the x pass has a kernel of its own");
    const int x = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int z = blockIdx.z + zoff;
    float vmax = 0.0f;
    if (x < nx && y < ny) {
        const size_t idx = ((size_t) z * ny + (size_t) y) * nx + (size_t) x;
        const int p = AXIS == 1 ? y : z;
        const size_t stride = AXIS == 1 ? (size_t) nx : (size_t) nx * ny;
        const float *__restrict__ q = src + (idx - (size_t) p * stride) + (size_t) A.base[p] * stride;
        const float *__restrict__ w = A.wT + p;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = 0; j < A.B; j++) {
            const float wj = __ldg(w + (size_t) j * A.n);
            const float4 v = *reinterpret_cast<const float4 *>(q + (size_t) j * stride);
            a.x = fmaf(wj, v.x, a.x);
            a.y = fmaf(wj, v.y, a.y);
            a.z = fmaf(wj, v.z, a.z);
            a.w = fmaf(wj, v.w, a.w);
        }
        *reinterpret_cast<float4 *>(dst + idx) = a;
        if (EPI) {
            if (dogmax != nullptr) {
                const float4 pq = *reinterpret_cast<const float4 *>(dog_prev + idx);
                vmax = fmaxf(fmaxf(fabsf(__fsub_rn(pq.x, a.x)), fabsf(__fsub_rn(pq.y, a.y))),
                             fmaxf(fabsf(__fsub_rn(pq.z, a.z)), fabsf(__fsub_rn(pq.w, a.w))));
            }
            if (ds_dst != nullptr && !((y | z) & 1)) {
                const int dnx = nx >> 1, dny = ny >> 1, dnz = nz >> 1;
                const int qx = x >> 1, qy = y >> 1, qz = z >> 1;
                if (qy < dny && qz < dnz) {
                    float *o = ds_dst + ((size_t) qz * dny + qy) * dnx + qx;
                    if (qx < dnx) o[0] = a.x;
                    if (qx + 1 < dnx) o[1] = a.z;
                }
            }
        }
    }
    if (EPI) {
        if (dogmax != nullptr) block_atomic_max(vmax, dogmax);
    }
}

}  // namespace s3d
