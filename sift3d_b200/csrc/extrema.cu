// extrema.cu -- DoG extremum detection fused with the DoG subtraction, and the
// deterministic ordering of the candidate list.
//
// Reference semantics (sift3d/sift.c:735-871): for s in {0..num_kp_levels-1}, voxels in
// [1, n-2]^3, D_s = G_s - G_{s+1} (build_dog, sift.c:713-732); accept iff
// |v| > (float)(peak_thresh * dogmax_s) and v is strictly greater (or strictly smaller) than
// its 6 face neighbours in D_s and the same voxel in D_{s-1} and D_{s+1}
// (CMP_PREV/CMP_CUR/CMP_NEXT, sift.c:798-811; CUBOID_EXTREMA is off, sift.c:24).
// The DoG pyramid is never written to HBM: every D value is one fp32 subtraction of two
// Gaussian levels, which is exactly what im_subtract stores (imutil.c:734-737).
#include "common.cuh"

namespace s3d {

struct ExtremaParams {
    const float *g[SIFT3D_CUDA_MAX_FILTERS + 1];  // Gaussian levels s = -1 .. kp+1
    float thresh[SIFT3D_CUDA_MAX_FILTERS];        // per keypoint level
    int num_kp_levels;
    int nx, ny, nz;
    int z0, z1;  // planes [z0, z1) are scanned (z-slab mode: the planes this rank owns)
    int s_off;   // keypoint level of g[1] (a launch may cover a sub-range of the levels)
};

__global__ void __launch_bounds__(256)
extrema_kernel(const ExtremaParams P, const float *__restrict__ dogmax, float peak_thresh_f,
               double peak_thresh, Cand *__restrict__ cand, int cap, int *__restrict__ count) {
    const int nx = P.nx, ny = P.ny, nz = P.nz;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    const int z = blockIdx.z + P.z0;
    const bool inside = x >= 1 && x <= nx - 2 && y >= 1 && y <= ny - 2 && z >= 1 && z <= nz - 2 && z < P.z1;
    const size_t ys = (size_t) nx, zs = (size_t) nx * ny;
    const size_t idx = (size_t) z * zs + (size_t) y * ys + x;
    const unsigned lane = threadIdx.x & 31u;

    for (int s = 0; s < P.num_kp_levels; s++) {
        bool hit = false;
        float v = 0.0f;
        if (inside) {
            // DoG level s lives between Gaussian levels s and s+1, i.e. g[s+1], g[s+2].
            const float *__restrict__ ga = P.g[s + 1];
            const float *__restrict__ gb = P.g[s + 2];
            v = __fsub_rn(ga[idx], gb[idx]);
            // peak_thresh = sift3d->peak_thresh * dogmax, stored to a float (sift.c:829)
            const float thr = (float) (peak_thresh * (double) dogmax[s]);
            if (v > thr || v < -thr) {
                const float prev = __fsub_rn(P.g[s][idx], ga[idx]);
                const float next = __fsub_rn(gb[idx], P.g[s + 3][idx]);
#define DV(off) __fsub_rn(ga[idx + (off)], gb[idx + (off)])
                const float xp = DV(1), xm = __fsub_rn(ga[idx - 1], gb[idx - 1]);
                const float yp = DV(ys), ym = __fsub_rn(ga[idx - ys], gb[idx - ys]);
                const float zp = DV(zs), zm = __fsub_rn(ga[idx - zs], gb[idx - zs]);
#undef DV
                const bool is_max = v > prev && v > xp && v > xm && v > yp && v > ym && v > zm &&
                                    v > zp && v > next;
                const bool is_min = v < prev && v < xp && v < xm && v < yp && v < ym && v < zm &&
                                    v < zp && v < next;
                hit = is_max || is_min;
            }
        }
        // warp-aggregated append
        const unsigned ballot = __ballot_sync(0xffffffffu, hit);
        if (ballot) {
            int base = 0;
            const int leader = __ffs(ballot) - 1;
            if ((int) lane == leader) base = atomicAdd(count, __popc(ballot));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (hit) {
                const int pos = base + __popc(ballot & ((1u << lane) - 1u));
                if (pos < cap) {
                    Cand c;
                    c.key = ((unsigned long long) (s + P.s_off) << 60) | ((unsigned long long) z << 40) |
                            ((unsigned long long) y << 20) | (unsigned long long) x;
                    c.strength = fabsf(v);
                    c.pad = 0;
                    cand[pos] = c;
                }
            }
        }
    }
    (void) peak_thresh_f;
}

// Vectorised variant for rows that are float4 addressable (nx % 4 == 0): a thread owns 4
// consecutive voxels, loads the kp+1 Gaussian levels that bound the keypoint DoG levels once
// (independent LDG.128s in flight together) and forms all DoG values in registers. Only
// 10-20 % of the (voxel, level) pairs pass the contrast threshold, scattered over the lanes:
// checking their 8 neighbours in place leaves most lanes idle (ncu: 11.7 of 32 threads active
// per instruction). So each warp first compacts its passing pairs into a shared-memory list
// (their final sort keys), then walks the list 32 pairs at a time with every lane busy.
template <int KP>
__global__ void __launch_bounds__(256)
extrema_vec4_kernel(const ExtremaParams P, const float *__restrict__ dogmax, double peak_thresh,
                    Cand *__restrict__ cand, int cap, int *__restrict__ count) {
    constexpr int kp = KP;
    constexpr int kItems = 32 * 4 * KP;  // worst case: every pair of the warp passes
    __shared__ unsigned long long items[8][kItems];
    const int nx = P.nx, ny = P.ny, nz = P.nz;
    const int nx4 = nx >> 2;
    const int x4 = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int z = blockIdx.z + P.z0;
    const bool row_ok = x4 < nx4 && y >= 1 && y <= ny - 2 && z >= 1 && z <= nz - 2 && z < P.z1;
    const size_t ys = (size_t) nx, zs = (size_t) nx * ny;
    const size_t idx0 = (size_t) z * zs + (size_t) y * ys + (size_t) x4 * 4;
    const unsigned tid = threadIdx.y * blockDim.x + threadIdx.x;
    const unsigned lane = tid & 31u, warp = tid >> 5;
    const unsigned full = 0xffffffffu;

    // ---- contrast test of this thread's 4 x kp pairs
    unsigned pass = 0;  // bit s*4 + c
    if (row_ok) {
        float4 g[KP + 1];  // levels 1 .. kp+1 of P.g (DoG s = g[s+1] - g[s+2])
#pragma unroll
        for (int l = 0; l <= kp; l++) g[l] = *reinterpret_cast<const float4 *>(P.g[l + 1] + idx0);
#pragma unroll
        for (int s = 0; s < kp; s++) {
            const float thr = (float) (peak_thresh * (double) dogmax[s]);
            const float4 a = g[s], b = g[s + 1];
            const float v4[4] = {__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z),
                                 __fsub_rn(a.w, b.w)};
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int x = x4 * 4 + c;
                if ((v4[c] > thr || v4[c] < -thr) && x >= 1 && x <= nx - 2) pass |= 1u << (s * 4 + c);
            }
        }
    }
    // ---- compact the passing pairs of the warp
    const int mine = __popc(pass);
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(full, incl, o);
        if ((int) lane >= o) incl += v;
    }
    const int total = __shfl_sync(full, incl, 31);
    if (total == 0) return;  // warp-uniform
    {
        int pos = incl - mine;
        unsigned m = pass;
        while (m) {
            const int bit = __ffs(m) - 1;
            m &= m - 1;
            const int s = bit >> 2, c = bit & 3;
            items[warp][pos++] = ((unsigned long long) (s + P.s_off) << 60) | ((unsigned long long) z << 40) |
                                 ((unsigned long long) y << 20) | (unsigned long long) (x4 * 4 + c);
        }
    }
    __syncwarp();
    // ---- 8-neighbour test (sift.c:798-811), one pair per lane
    for (int i0 = 0; i0 < total; i0 += 32) {
        const int i = i0 + (int) lane;
        bool hit = false;
        unsigned long long key = 0;
        float v = 0.0f;
        if (i < total) {
            key = items[warp][i];
            const int s = (int) (key >> 60) - P.s_off;
            const size_t idx = (size_t) ((key >> 40) & 0xFFFFF) * zs + (size_t) ((key >> 20) & 0xFFFFF) * ys +
                               (size_t) (key & 0xFFFFF);
            const float *__restrict__ ga = P.g[s + 1];
            const float *__restrict__ gb = P.g[s + 2];
            v = __fsub_rn(ga[idx], gb[idx]);
            const float prev = __fsub_rn(P.g[s][idx], ga[idx]);
            const float next = __fsub_rn(gb[idx], P.g[s + 3][idx]);
            const float xp = __fsub_rn(ga[idx + 1], gb[idx + 1]), xm = __fsub_rn(ga[idx - 1], gb[idx - 1]);
            const float yp = __fsub_rn(ga[idx + ys], gb[idx + ys]), ym = __fsub_rn(ga[idx - ys], gb[idx - ys]);
            const float zp = __fsub_rn(ga[idx + zs], gb[idx + zs]), zm = __fsub_rn(ga[idx - zs], gb[idx - zs]);
            const bool is_max = v > prev && v > xp && v > xm && v > yp && v > ym && v > zm && v > zp &&
                                v > next;
            const bool is_min = v < prev && v < xp && v < xm && v < yp && v < ym && v < zm && v < zp &&
                                v < next;
            hit = is_max || is_min;
        }
        const unsigned ballot = __ballot_sync(full, hit);
        if (ballot) {
            int base = 0;
            const int leader = __ffs(ballot) - 1;
            if ((int) lane == leader) base = atomicAdd(count, __popc(ballot));
            base = __shfl_sync(full, base, leader);
            if (hit) {
                const int pos = base + __popc(ballot & ((1u << lane) - 1u));
                if (pos < cap) {
                    Cand cd;
                    cd.key = key;
                    cd.strength = fabsf(v);
                    cd.pad = 0;
                    cand[pos] = cd;
                }
            }
        }
    }
}

// Rank sort: candidate i goes to slot #{j : key_j < key_i}. Keys are unique (one per
// voxel and level) so the ranks are a permutation; the result is the reference's serial
// scan order (s, z, y, x) regardless of the order the atomics appended them in.
// One launch orders every octave: blockIdx.y = octave, counts are read on the device.
struct OrderParams {
    int base[SIFT3D_CUDA_MAX_OCTAVES];  // first slot of the octave's candidate region
    int cap[SIFT3D_CUDA_MAX_OCTAVES];
};

// Regions with at most kOrderSmall candidates are rank-sorted directly (O(n^2 / 256) per CTA:
// a few microseconds for the few thousand candidates of a 256^3 octave). Busier regions (1024^3:
// 300k candidates in octave 0, 130k in a 128^3 octave of noise-free blobs -- 10 ms quadratic) are
// first grouped by their (level, plane) bucket and rank-sorted inside each bucket. Both sets
// of kernels are always enqueued and choose on the device by the counted number, so the
// schedule needs no host round trip.
constexpr int kOrderSmall = 6144;

__global__ void __launch_bounds__(256)
order_all_kernel(const Cand *__restrict__ raw, Cand *__restrict__ sorted, const OrderParams P,
                 const int *__restrict__ counts, int o_begin) {
    __shared__ unsigned long long tile[256];
    const int o = blockIdx.y + o_begin;
    const int n = min(counts[o], P.cap[o]);
    if (n > kOrderSmall) return;  // the bucket kernels take it
    const Cand *__restrict__ in = raw + P.base[o];
    Cand *__restrict__ out = sorted + P.base[o];
    for (int first = blockIdx.x * 256; first < n; first += gridDim.x * 256) {
        const int i = first + threadIdx.x;
        Cand mine;
        mine.key = ~0ull;
        mine.strength = 0.f;
        mine.pad = 0;
        if (i < n) mine = in[i];
        int rank = 0;
        for (int base = 0; base < n; base += 256) {
            const int j = base + threadIdx.x;
            __syncthreads();
            tile[threadIdx.x] = j < n ? in[j].key : ~0ull;
            __syncthreads();
            const int lim = min(256, n - base);
#pragma unroll 8
            for (int k = 0; k < lim; k++) rank += tile[k] < mine.key ? 1 : 0;
        }
        if (i < n) out[rank] = mine;
    }
}

// ---- bucket path: bucket = level * nz + plane (the two leading fields of the key)
struct BucketParams {
    int base, cap;     // candidate region of the octave
    int nz, nbkt;      // planes of the octave, buckets = keypoint levels * nz
    int *bcount;       // [nbkt] members per bucket (zeroed before the launch)
    int *bstart;       // [nbkt] first slot of the bucket in the grouped / sorted region
    int *cursor;       // [nbkt] scatter cursors
};
__device__ __forceinline__ int bucket_of(unsigned long long key, int nz) {
    return (int) (key >> 60) * nz + (int) ((key >> 40) & 0xfffffull);
}

__global__ void __launch_bounds__(256)
bucket_count_kernel(const Cand *__restrict__ raw, const BucketParams B, const int *__restrict__ count) {
    const int n = min(*count, B.cap);
    if (n <= kOrderSmall) return;
    const Cand *__restrict__ in = raw + B.base;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256)
        atomicAdd(&B.bcount[bucket_of(in[i].key, B.nz)], 1);
}

__global__ void __launch_bounds__(1024)
bucket_scan_kernel(const BucketParams B, const int *__restrict__ count) {
    __shared__ int wsum[32];
    __shared__ int carry;
    const int n = min(*count, B.cap);
    if (n <= kOrderSmall) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int first = 0; first < B.nbkt; first += 1024) {
        const int b = first + tid;
        const int v = b < B.nbkt ? B.bcount[b] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = wsum[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            wsum[lane] = wi - w;  // exclusive over the warps
        }
        __syncthreads();
        const int excl = carry + wsum[warp] + incl - v;
        if (b < B.nbkt) {
            B.bstart[b] = excl;
            B.cursor[b] = 0;
        }
        __syncthreads();
        if (tid == 1023) carry = excl + v;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
bucket_scatter_kernel(const Cand *__restrict__ raw, Cand *__restrict__ grouped, const BucketParams B,
                      const int *__restrict__ count) {
    const int n = min(*count, B.cap);
    if (n <= kOrderSmall) return;
    const Cand *__restrict__ in = raw + B.base;
    Cand *__restrict__ out = grouped + B.base;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
        const Cand c = in[i];
        const int b = bucket_of(c.key, B.nz);
        out[B.bstart[b] + atomicAdd(&B.cursor[b], 1)] = c;
    }
}

// one CTA per bucket (grid-stride over the buckets): rank sort of its members by the whole key
__global__ void __launch_bounds__(128)
bucket_rank_kernel(const Cand *__restrict__ grouped, Cand *__restrict__ sorted, const BucketParams B,
                   const int *__restrict__ count) {
    __shared__ unsigned long long tile[128];
    const int n = min(*count, B.cap);
    if (n <= kOrderSmall) return;
    for (int b = blockIdx.x; b < B.nbkt; b += gridDim.x) {
        const int m = B.bcount[b];
        if (m == 0) continue;
        const Cand *__restrict__ in = grouped + B.base + B.bstart[b];
        Cand *__restrict__ out = sorted + B.base + B.bstart[b];
        for (int first = 0; first < m; first += 128) {
            const int i = first + threadIdx.x;
            Cand mine;
            mine.key = ~0ull;
            mine.strength = 0.f;
            mine.pad = 0;
            if (i < m) mine = in[i];
            int rank = 0;
            for (int base = 0; base < m; base += 128) {
                const int j = base + threadIdx.x;
                __syncthreads();
                tile[threadIdx.x] = j < m ? in[j].key : ~0ull;
                __syncthreads();
                const int lim = min(128, m - base);
                for (int k = 0; k < lim; k++) rank += tile[k] < mine.key ? 1 : 0;
            }
            if (i < m) out[rank] = mine;
        }
        __syncthreads();
    }
}

int launch_extrema(cudaStream_t st, const float *const *g, int num_kp_levels, int nx, int ny,
                   int nz, const float *d_dogmax, double peak_thresh, Cand *d_cand, int cap,
                   int *d_count, uint64_t *launches, int z0, int z1, int s_begin, int s_end) {
    if (z1 < 0) {
        z0 = 0;
        z1 = nz;
    }
    // keypoint levels [s_begin, s_end) only: level s needs g[s] .. g[s+3] and dogmax[s]
    if (s_end < 0) s_end = num_kp_levels;
    if (s_end <= s_begin) return 0;
    g += s_begin;
    d_dogmax += s_begin;
    num_kp_levels = s_end - s_begin;
    ExtremaParams P;
    for (int i = 0; i < num_kp_levels + 3; i++) P.g[i] = g[i];
    P.s_off = s_begin;
    P.num_kp_levels = num_kp_levels;
    P.nx = nx;
    P.ny = ny;
    P.nz = nz;
    P.z0 = z0;
    P.z1 = z1;
    bool aligned = nx % 4 == 0;
    for (int i = 0; i < num_kp_levels + 3; i++) aligned = aligned && ((uintptr_t) g[i] & 15) == 0;
    if (aligned && num_kp_levels >= 1 && num_kp_levels <= 4) {
        const int nx4 = nx / 4;
        int bx = 1;
        while (bx < nx4 && bx < 64) bx <<= 1;  // power of two <= 64 so that warps stay row aligned
        dim3 block((unsigned) bx, (unsigned) (256 / bx), 1);
        dim3 grid((unsigned) ((nx4 + bx - 1) / bx), (unsigned) ((ny + block.y - 1) / block.y), (unsigned) (z1 - z0));
        switch (num_kp_levels) {
        case 1: extrema_vec4_kernel<1><<<grid, block, 0, st>>>(P, d_dogmax, peak_thresh, d_cand, cap, d_count); break;
        case 2: extrema_vec4_kernel<2><<<grid, block, 0, st>>>(P, d_dogmax, peak_thresh, d_cand, cap, d_count); break;
        case 3: extrema_vec4_kernel<3><<<grid, block, 0, st>>>(P, d_dogmax, peak_thresh, d_cand, cap, d_count); break;
        default: extrema_vec4_kernel<4><<<grid, block, 0, st>>>(P, d_dogmax, peak_thresh, d_cand, cap, d_count); break;
        }
    } else {
        dim3 block(128, 1, 1);
        dim3 grid((unsigned) ((nx + 127) / 128), (unsigned) ny, (unsigned) (z1 - z0));
        extrema_kernel<<<grid, block, 0, st>>>(P, d_dogmax, 0.f, peak_thresh, d_cand, cap, d_count);
    }
    if (launches) *launches += 1;
    S3D_CK(cudaGetLastError());
    return 0;
}

int launch_order_all(cudaStream_t st, const Cand *d_raw, Cand *d_sorted, const int *base,
                     const int *cap, int noct, const int *d_counts, uint64_t *launches,
                     int o_begin, int o_end, const OrderScratch *scratch) {
    OrderParams P;
    for (int o = 0; o < SIFT3D_CUDA_MAX_OCTAVES; o++) {
        P.base[o] = o < noct ? base[o] : 0;
        P.cap[o] = o < noct ? cap[o] : 0;
    }
    if (o_end < 0) o_end = noct;
    if (o_end <= o_begin) return 0;
    order_all_kernel<<<dim3(32u, (unsigned) (o_end - o_begin)), 256, 0, st>>>(d_raw, d_sorted, P, d_counts,
                                                                             o_begin);
    if (launches) *launches += 1;
    // regions that can hold more than the direct sort takes also get the bucket kernels
    for (int o = o_begin; o < o_end && o < noct && scratch != nullptr; o++) {
        if (cap[o] <= kOrderSmall || scratch->grouped == nullptr) continue;
        BucketParams B;
        B.base = base[o];
        B.cap = cap[o];
        B.nz = scratch->nz[o];
        B.nbkt = scratch->nbkt[o];
        B.bcount = scratch->ints + scratch->off[o];
        B.bstart = B.bcount + B.nbkt;
        B.cursor = B.bstart + B.nbkt;
        S3D_CK(cudaMemsetAsync(B.bcount, 0, sizeof(int) * (size_t) B.nbkt, st));
        const unsigned g = (unsigned) kNumSMs * 4;
        bucket_count_kernel<<<g, 256, 0, st>>>(d_raw, B, d_counts + o);
        bucket_scan_kernel<<<1, 1024, 0, st>>>(B, d_counts + o);
        bucket_scatter_kernel<<<g, 256, 0, st>>>(d_raw, scratch->grouped, B, d_counts + o);
        bucket_rank_kernel<<<(unsigned) (B.nbkt < kNumSMs * 16 ? B.nbkt : kNumSMs * 16), 128, 0, st>>>(
            scratch->grouped, d_sorted, B, d_counts + o);
        if (launches) *launches += 4;
    }
    S3D_CK(cudaGetLastError());
    return 0;
}

}  // namespace s3d
