// multi.cu -- one tall volume over several GPUs of one process:
// z-slab sharding of the detect + describe path with halo exchange over NVLink peer copies.
//
// The reference has no multi-device path; this is the decomposition of its octave loops
// (build_gpyr sift.c:662-711, detect_extrema sift.c:735-871, assign_orientations
// sift.c:1109-1167, extract_descrip sift.c:1442-1536) over z, arranged so that every device
// computes exactly the values one device would (sift3d_cuda_detect_slab):
//   * device r owns the planes [g0, g1) (multiples of 2^K, K = sharded octaves) and holds the
//     array [a0, a1) = owned planes + `halo` planes on each cut face;
//   * one host thread per device runs the slab detect; its callbacks never drain a stream. After
//     a level is produced the thread records an event and ENQUEUES, behind the neighbours'
//     events, peer copies (cudaMemcpyPeerAsync) of just the planes the level's consumers read:
//     the reach of the next blur for every level, and once per keypoint level the planes the
//     orientation / descriptor windows reach;
//   * max|image| and max|DoG| are max-reduced in-stream (peer copies of the scalars + a kernel;
//     max is exact and idempotent);
//   * octaves >= K are too small to shard: their seed level is gathered on device 0, which runs
//     them through sift3d_cuda_detect_from_seed;
//   * every device describes its own keypoints; the host library merges the per-device lists in
//     the reference's scan order (o, s, z, y, x) = (o, s, device, local order).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <new>
#include <thread>
#include <vector>

#include "common.cuh"

using namespace s3d;

namespace {

constexpr int kMaxRanks = SIFT3D_CUDA_MULTI_MAX_DEVICES;
constexpr int kRing = 512;  // events per rank: more than the callbacks of one detect

__global__ void rows_max_kernel(const float *__restrict__ rows, int nrows, int stride, int count,
                                float *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float m = out[i];
    for (int q = 0; q < nrows; q++) m = fmaxf(m, rows[(size_t) q * stride + i]);
    out[i] = m;
}

}  // namespace

struct sift3d_cuda_multi {
    sift3d_cuda_multi_desc D;
    int P = 0, K = 0, kp = 0;
    std::vector<sift3d_cuda_plan_desc> plans;  // per rank
    sift3d_cuda_plan_desc tail_plan;
    bool has_tail = false;
    sift3d_cuda_ctx *ctx[kMaxRanks] = {};
    sift3d_cuda_ctx *tail = nullptr;
    cudaStream_t stream[kMaxRanks] = {};
    sift3d_cuda_slab_desc slab[kMaxRanks];
    sift3d_cuda_slab_callbacks cb[kMaxRanks];
    struct Self { sift3d_cuda_multi *m; int r; } self[kMaxRanks];
    // in-stream synchronisation: callback number k of rank r records ev[r][k % kRing]
    std::vector<cudaEvent_t> ev[kMaxRanks];
    std::atomic<int> published[kMaxRanks];
    int step[kMaxRanks] = {};
    float *vals_ptr[kMaxRanks] = {};   // this step's reduce operand of every rank
    float *gather[kMaxRanks] = {};     // [P][gather_cap] staging rows on every device
    int gather_cap = 0;
    std::atomic<int> failed{0};
    float *seed = nullptr;             // level (K, -1) of the whole volume, on device of rank 0
    // results, per rank (+ one slot for the tail octaves)
    std::vector<sift3d_cuda_keypoint> cand[kMaxRanks + 1];
    sift3d_cuda_multi_timings tm;
    size_t exchanged_bytes[kMaxRanks] = {};
};

namespace {

int held0(const sift3d_cuda_multi *m, int r, int o) { return m->D.a0[r] >> o; }

// Planes (octave-local, each side) of level (o, s) that its consumers on a neighbouring slab read
int planes_needed(const sift3d_cuda_multi *m, int o, int s) {
    const sift3d_cuda_plan_desc &pl = m->plans[0];
    const double uz = ldexp(m->D.units[2], o);
    int need = 2;  // extrema: z +- 1 of every Gaussian level
    const int nfilt = m->kp + 2;
    if (s + 1 < nfilt) {  // the blur that produces level s+1 (filter index s+1) reads level s
        const int hw = pl.octave[s + 1].width / 2;
        const int uhw = (int) ceilf((float) hw * (float) (1.0 / uz));
        need = uhw + 2 > need ? uhw + 2 : need;
    }
    if (s >= 0 && s < m->kp) need = m->D.window_planes[o] > need ? m->D.window_planes[o] : need;
    return need;
}

int cb_exchange(void *user, int o, int s, float *d_level, int nx, int ny, int nz, int own0, int own1) {
    sift3d_cuda_multi::Self *me = (sift3d_cuda_multi::Self *) user;
    sift3d_cuda_multi *m = me->m;
    const int r = me->r, P = m->P;
    (void) nz; (void) own0; (void) own1;
    const int k = ++m->step[r];
    S3D_CK(cudaEventRecord(m->ev[r][k % kRing], m->stream[r]));
    m->published[r].store(k, std::memory_order_release);
    if (o >= m->K) return 0;  // the seed of the unsharded octaves is gathered whole afterwards
    const int need = planes_needed(m, o, s);
    const size_t plane = (size_t) nx * ny;
    const int nzo = m->D.nz >> o;
    const int base = held0(m, r, o);
    const int og0 = m->D.g0[r] >> o, og1 = m->D.g1[r] >> o;
    // wanted global planes: [og0 - need, og0) and [og1, og1 + need), inside the held array
    const int h0 = m->D.a0[r] >> o, h1 = (m->D.a1[r] >> o) < nzo ? (m->D.a1[r] >> o) : nzo;
    const int lo0 = og0 - need > h0 ? og0 - need : h0, lo1 = og0;
    const int hi0 = og1, hi1 = og1 + need < h1 ? og1 + need : h1;
    for (int q = 0; q < P; q++) {
        if (q == r) continue;
        const int q0 = m->D.g0[q] >> o, q1 = m->D.g1[q] >> o;
        const int ranges[2][2] = {{lo0 > q0 ? lo0 : q0, lo1 < q1 ? lo1 : q1},
                                  {hi0 > q0 ? hi0 : q0, hi1 < q1 ? hi1 : q1}};
        bool waited = false;
        for (int t = 0; t < 2; t++) {
            const int p0 = ranges[t][0], p1 = ranges[t][1];
            if (p0 >= p1) continue;
            if (!waited) {
                while (m->published[q].load(std::memory_order_acquire) < k) {
                    if (m->failed.load()) return -1;
                    std::this_thread::yield();
                }
                S3D_CK(cudaStreamWaitEvent(m->stream[r], m->ev[q][k % kRing], 0));
                waited = true;
            }
            const float *src = sift3d_cuda_level_ptr(m->ctx[q], o, s);
            if (src == nullptr) return -1;
            src += (size_t) (p0 - held0(m, q, o)) * plane;
            float *dst = d_level + (size_t) (p0 - base) * plane;
            const size_t bytes = (size_t) (p1 - p0) * plane * sizeof(float);
            S3D_CK(cudaMemcpyPeerAsync(dst, m->D.devices[r], src, m->D.devices[q], bytes, m->stream[r]));
            m->exchanged_bytes[r] += bytes;
        }
    }
    return 0;
}

int cb_reduce_max(void *user, float *d_vals, int count) {
    sift3d_cuda_multi::Self *me = (sift3d_cuda_multi::Self *) user;
    sift3d_cuda_multi *m = me->m;
    const int r = me->r, P = m->P;
    if (count > m->gather_cap) {
        S3D_ERR("sift3d_cuda_multi: reduce of %d values exceeds the staging rows (%d)\n", count, m->gather_cap);
        return -1;
    }
    const int k = ++m->step[r];
    m->vals_ptr[r] = d_vals;
    S3D_CK(cudaEventRecord(m->ev[r][k % kRing], m->stream[r]));
    m->published[r].store(k, std::memory_order_release);
    if (P == 1) return 0;
    int rows = 0;
    for (int q = 0; q < P; q++) {
        if (q == r) continue;
        while (m->published[q].load(std::memory_order_acquire) < k) {
            if (m->failed.load()) return -1;
            std::this_thread::yield();
        }
        S3D_CK(cudaStreamWaitEvent(m->stream[r], m->ev[q][k % kRing], 0));
        // (a peer may already have replaced its values by the global maximum: max is idempotent)
        S3D_CK(cudaMemcpyPeerAsync(m->gather[r] + (size_t) rows * m->gather_cap, m->D.devices[r],
                                   m->vals_ptr[q], m->D.devices[q], (size_t) count * sizeof(float),
                                   m->stream[r]));
        rows++;
    }
    rows_max_kernel<<<(count + 127) / 128, 128, 0, m->stream[r]>>>(m->gather[r], rows, m->gather_cap, count, d_vals);
    S3D_CK(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

void sift3d_cuda_multi_destroy(sift3d_cuda_multi *m) {
    if (!m) return;
    for (int r = 0; r < m->P; r++) {
        if (m->ctx[r]) {
            cudaSetDevice(m->D.devices[r]);
            sift3d_cuda_synchronize(m->ctx[r]);
        }
    }
    for (int r = 0; r < m->P; r++) {
        cudaSetDevice(m->D.devices[r]);
        for (cudaEvent_t e : m->ev[r]) cudaEventDestroy(e);
        if (m->gather[r]) cudaFree(m->gather[r]);
        if (m->ctx[r]) sift3d_cuda_ctx_destroy(m->ctx[r]);
    }
    if (m->P > 0) cudaSetDevice(m->D.devices[0]);
    if (m->seed) cudaFree(m->seed);
    if (m->tail) sift3d_cuda_ctx_destroy(m->tail);
    delete m;
}

sift3d_cuda_multi *sift3d_cuda_multi_create(const sift3d_cuda_multi_desc *d,
                                            const sift3d_cuda_plan_desc *slab_plans,
                                            const sift3d_cuda_plan_desc *tail_plan) {
    if (!d || !slab_plans || d->num_devices < 1 || d->num_devices > kMaxRanks ||
        d->sharded_octaves < 1 || d->sharded_octaves >= SIFT3D_CUDA_MAX_OCTAVES) {
        S3D_ERR("sift3d_cuda_multi_create: invalid description\n");
        return nullptr;
    }
    sift3d_cuda_multi *m = new (std::nothrow) sift3d_cuda_multi();
    if (!m) return nullptr;
    m->D = *d;
    m->P = d->num_devices;
    m->K = d->sharded_octaves;
    m->kp = slab_plans[0].num_kp_levels;
    m->plans.assign(slab_plans, slab_plans + m->P);
    m->has_tail = tail_plan != nullptr;
    if (tail_plan) m->tail_plan = *tail_plan;
    memset(&m->tm, 0, sizeof(m->tm));
    const int ndev = sift3d_cuda_device_count();
    bool ok = true;
    for (int r = 0; r < m->P && ok; r++) {
        m->published[r].store(0);
        const int dev = d->devices[r];
        const sift3d_cuda_plan_desc &pl = m->plans[r];
        if (dev < 0 || dev >= ndev || d->g0[r] < d->a0[r] || d->g1[r] > d->a1[r] || d->g0[r] >= d->g1[r] ||
            pl.nz != d->a1[r] - d->a0[r] || pl.nx != d->nx || pl.ny != d->ny ||
            pl.num_octaves < m->K || pl.num_octaves > m->K + 1) {
            S3D_ERR("sift3d_cuda_multi_create: bad slab %d\n", r);
            ok = false;
            break;
        }
        m->ctx[r] = sift3d_cuda_ctx_create(dev);
        if (!m->ctx[r] || sift3d_cuda_plan(m->ctx[r], &pl)) {
            ok = false;
            break;
        }
        m->stream[r] = (cudaStream_t) sift3d_cuda_ctx_stream(m->ctx[r]);
        ok = ok && cudaSetDevice(dev) == cudaSuccess;
        m->ev[r].assign(kRing, nullptr);
        for (cudaEvent_t &e : m->ev[r]) ok = ok && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
        m->gather_cap = 1 + SIFT3D_CUDA_MAX_OCTAVES * SIFT3D_CUDA_MAX_FILTERS;
        ok = ok && cudaMalloc((void **) &m->gather[r], (size_t) m->P * m->gather_cap * sizeof(float)) == cudaSuccess;
        // direct peer copies where the hardware allows (else the driver stages them)
        for (int q = 0; q < m->P && ok; q++) {
            const int pd = d->devices[q];
            int can = 0;
            if (pd != dev && cudaDeviceCanAccessPeer(&can, dev, pd) == cudaSuccess && can) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(pd, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = false;
                cudaGetLastError();
            }
        }
        sift3d_cuda_slab_desc &S = m->slab[r];
        memset(&S, 0, sizeof(S));
        S.num_sharded_octaves = m->K;
        for (int o = 0; o < pl.num_octaves; o++) {
            S.own0[o] = (d->g0[r] >> o) - (d->a0[r] >> o);
            S.own1[o] = (d->g1[r] >> o) - (d->a0[r] >> o);
            S.org[o] = d->a0[r] >> o;
            S.vol_nz[o] = d->nz >> o;
        }
        m->self[r].m = m;
        m->self[r].r = r;
        m->cb[r].user = &m->self[r];
        m->cb[r].reduce_max = cb_reduce_max;
        m->cb[r].exchange = cb_exchange;
        m->cb[r].in_stream = 1;
        m->cb[r].input_reach = pl.first.width / 2 + 2;
    }
    if (ok && m->has_tail) {
        const int dev0 = d->devices[0];
        m->tail = sift3d_cuda_ctx_create(dev0);
        ok = m->tail && sift3d_cuda_plan(m->tail, &m->tail_plan) == 0 &&
             sift3d_cuda_set_base_octave(m->tail, m->K) == 0 && cudaSetDevice(dev0) == cudaSuccess &&
             cudaMalloc((void **) &m->seed, (size_t) m->tail_plan.nx * m->tail_plan.ny * m->tail_plan.nz *
                                                sizeof(float)) == cudaSuccess;
        if (ok && (m->tail_plan.nx != d->nx >> m->K || m->tail_plan.ny != d->ny >> m->K ||
                   m->tail_plan.nz != d->nz >> m->K)) {
            S3D_ERR("sift3d_cuda_multi_create: tail plan does not match the volume\n");
            ok = false;
        }
    }
    if (!ok) {
        S3D_ERR("sift3d_cuda_multi_create failed: %s\n", cudaGetErrorString(cudaGetLastError()));
        sift3d_cuda_multi_destroy(m);
        return nullptr;
    }
    return m;
}

}  // extern "C"

namespace {

double now_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return 1e3 * (double) ts.tv_sec + 1e-6 * (double) ts.tv_nsec;
}

int fetch(sift3d_cuda_ctx *ctx, int n, std::vector<sift3d_cuda_keypoint> &cand) {
    cand.resize((size_t) n);
    if (n == 0) return 0;
    return sift3d_cuda_fetch_candidates(ctx, cand.data(), n) == n ? 0 : -1;
}

}  // namespace

extern "C" {

int sift3d_cuda_multi_detect(sift3d_cuda_multi *m, const float *host_image, double peak_thresh,
                             double corner_thresh) {
    if (!m || !host_image || m->failed.load()) {
        S3D_ERR("sift3d_cuda_multi_detect: unusable object or NULL image\n");
        return -1;
    }
    const int P = m->P, K = m->K;
    const size_t plane0 = (size_t) m->D.nx * m->D.ny;
    int status[kMaxRanks] = {};
    double t_detect[kMaxRanks] = {};
    const double t0 = now_ms();
    auto work = [&](int r) {
        int n = 0;
        const double ta = now_ms();
        int rc = sift3d_cuda_detect_slab(m->ctx[r], host_image + (size_t) m->D.a0[r] * plane0, 0, m->D.units,
                                         peak_thresh, corner_thresh, &m->slab[r], &m->cb[r], &n);
        t_detect[r] = now_ms() - ta;
        if (rc == 0) rc = fetch(m->ctx[r], n, m->cand[r]) ? -1 : 0;
        if (rc == 0)
            for (sift3d_cuda_keypoint &c : m->cand[r])  // slab-array z -> volume z (octave coordinates)
                c.z += (double) (m->D.a0[r] >> c.o);
        if (rc != 0 && rc != SIFT3D_CUDA_SLAB_RETRY) m->failed.store(1);
        status[r] = rc;
    };
    for (int attempt = 0;; attempt++) {
        {
            std::vector<std::thread> th;
            for (int r = 1; r < P; r++) th.emplace_back(work, r);
            work(0);
            for (std::thread &t : th) t.join();
        }
        bool retry = false;
        for (int r = 0; r < P; r++) {
            if (status[r] == SIFT3D_CUDA_SLAB_RETRY) {
                retry = true;
            } else if (status[r]) {
                S3D_ERR("sift3d_cuda_multi_detect: slab %d failed\n", r);
                return -1;
            }
        }
        if (!retry) break;
        // a busy image overflowed a candidate region (resized by now): every slab runs again
        if (attempt >= 2) {
            S3D_ERR("sift3d_cuda_multi_detect: candidate regions still too small\n");
            m->failed.store(1);
            return -1;
        }
        m->failed.store(0);
    }
    const double t1 = now_ms();
    m->cand[P].clear();
    if (m->has_tail) {
        // level (K, -1) of the whole volume from the owners' planes, then the small octaves
        const int dev0 = m->D.devices[0];
        S3D_CK(cudaSetDevice(dev0));
        const size_t plane = (size_t) (m->D.nx >> K) * (m->D.ny >> K);
        cudaStream_t s0 = m->stream[0];
        for (int r = 0; r < P; r++) {
            const int p0 = m->D.g0[r] >> K, p1 = m->D.g1[r] >> K;
            if (p0 >= p1) continue;
            const float *src = sift3d_cuda_level_ptr(m->ctx[r], K, -1);
            if (!src) return -1;
            src += (size_t) (p0 - (m->D.a0[r] >> K)) * plane;
            S3D_CK(cudaMemcpyPeerAsync(m->seed + (size_t) p0 * plane, dev0, src, m->D.devices[r],
                                       (size_t) (p1 - p0) * plane * sizeof(float), s0));
        }
        S3D_CK(cudaStreamSynchronize(s0));
        double tu[3] = {ldexp(m->D.units[0], K), ldexp(m->D.units[1], K), ldexp(m->D.units[2], K)};
        int n = 0;
        if (sift3d_cuda_detect_from_seed(m->tail, m->seed, 1, tu, peak_thresh, corner_thresh, &n) ||
            fetch(m->tail, n, m->cand[P])) {
            m->failed.store(1);
            return -1;
        }
    }
    const double t2 = now_ms();
    m->tm.slabs_ms = (float) (t1 - t0);
    m->tm.tail_ms = (float) (t2 - t1);
    m->tm.total_ms = (float) (t2 - t0);
    m->tm.max_detect_ms = m->tm.max_describe_ms = m->tm.describe_ms = 0.0f;
    for (int r = 0; r < P; r++)
        if (t_detect[r] > m->tm.max_detect_ms) m->tm.max_detect_ms = (float) t_detect[r];
    size_t ex = 0;
    for (int r = 0; r < P; r++) {
        ex += m->exchanged_bytes[r];
        m->exchanged_bytes[r] = 0;
    }
    m->tm.exchanged_mbytes = (float) ((double) ex / 1048576.0);
    return 0;
}

int sift3d_cuda_multi_results(sift3d_cuda_multi *m, int part, const sift3d_cuda_keypoint **cands) {
    if (!m || part < 0 || part > m->P) return -1;
    if (cands) *cands = m->cand[part].empty() ? nullptr : m->cand[part].data();
    return (int) m->cand[part].size();
}

int sift3d_cuda_multi_describe(sift3d_cuda_multi *m, float *desc, const int *const *rows) {
    if (!m || !desc || !rows || m->failed.load()) return -1;
    const int P = m->P;
    int status[kMaxRanks + 1] = {};
    double t_desc[kMaxRanks + 1] = {};
    const double t0 = now_ms();
    auto work = [&](int part) {
        const double ta = now_ms();
        sift3d_cuda_ctx *ctx = part < P ? m->ctx[part] : m->tail;
        std::vector<sift3d_cuda_keypoint> keys;
        keys.reserve(m->cand[part].size() / 3 + 16);
        for (const sift3d_cuda_keypoint &c : m->cand[part])
            if (c.valid) {
                keys.push_back(c);
                if (part < P) keys.back().z -= (double) (m->D.a0[part] >> c.o);  // back to slab-array z
            }
        int rc = 0;
        if (!keys.empty()) {
            if (rows[part] == nullptr) rc = -1;
            else rc = sift3d_cuda_describe_rows(ctx, keys.data(), (int) keys.size(), desc, 0, rows[part]);
        }
        status[part] = rc;
        t_desc[part] = now_ms() - ta;
    };
    {
        // the unsharded octaves live on device 0 next to slab 0: a thread of their own
        std::vector<std::thread> th;
        const int nparts = P + (m->has_tail ? 1 : 0);
        for (int part = 1; part < nparts; part++) th.emplace_back(work, part);
        work(0);
        for (std::thread &t : th) t.join();
    }
    for (int part = 0; part <= P; part++)
        if (status[part]) {
            S3D_ERR("sift3d_cuda_multi_describe: part %d failed\n", part);
            return -1;
        }
    m->tm.describe_ms = (float) (now_ms() - t0);
    for (int part = 0; part <= P; part++)
        if (t_desc[part] > m->tm.max_describe_ms) m->tm.max_describe_ms = (float) t_desc[part];
    m->tm.total_ms += m->tm.describe_ms;
    return 0;
}

int sift3d_cuda_multi_get_timings(sift3d_cuda_multi *m, sift3d_cuda_multi_timings *t) {
    if (!m || !t) return -1;
    *t = m->tm;
    return 0;
}

}  // extern "C"
