// sift3d_cuda.cu -- context, device arena, pipeline schedule and the C ABI of
// libsift3d_cuda.so (include/sift3d_cuda.h). One context = one device + one stream +
// one resident pyramid, mirroring "one detector keeps the last image's pyramid so that
// descriptors can be extracted later" (sift3d/sift.c:1622-1628).
#include <math.h>
#include <stdlib.h>

#include <atomic>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "common.cuh"

using namespace s3d;

namespace {

enum { EV_START, EV_H2D, EV_SCALE, EV_PYR, EV_EXTREMA, EV_ORDER, EV_ORIENT, EV_D2H, EV_COUNT };
enum { DV_START, DV_H2D, DV_KERNEL, DV_D2H, DV_COUNT };

struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
    int ensure(size_t need) {
        if (need <= bytes) return 0;
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
        S3D_CK(cudaMalloc(&p, need));
        bytes = need;
        return 0;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
    }
};

}  // namespace

struct sift3d_cuda_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[EV_COUNT] = {};
    cudaEvent_t dv[DV_COUNT] = {};
    uint64_t launches = 0;

    // plan
    bool planned = false;
    sift3d_cuda_plan_desc desc;
    int nlev = 0;  // Gaussian levels per octave = num_kp_levels + 3
    std::vector<Level> levels;          // [o * nlev + (s + 1)]
    DevBuf pyr, input, tmp_a, tmp_b;    // arena (tmp_*: scratch of the 3-pass blur, octave 0)
    DevBuf tmp_side;                    // scratch of octaves >= 1 (they run on their own streams)
    std::vector<size_t> side_off;       // [o]: offset of octave o's two scratch volumes in tmp_side
    // octave o >= 1 runs on oct_stream[o] as soon as its seed level exists (sift.c:694-705);
    // ev_seed[o] = level (o+1, -1) written, ev_done[o] = octave o finished (levels + extrema)
    cudaStream_t oct_stream[SIFT3D_CUDA_MAX_OCTAVES] = {};
    cudaEvent_t ev_seed[SIFT3D_CUDA_MAX_OCTAVES] = {};
    cudaEvent_t ev_done[SIFT3D_CUDA_MAX_OCTAVES] = {};
    bool serial = false;                // $SIFT3D_CUDA_SERIAL: everything on one stream
    bool exact = false;                 // 0-ULP blur arithmetic (sift3d_cuda_set_exact, $SIFT3D_CUDA_EXACT)
    // The launches of one detect (everything behind the first blur, whose source is the caller's
    // image) as a CUDA graph: captured the second time a detect runs with the same geometry,
    // units, thresholds and arithmetic mode -- the reference's own "re-plan only if the
    // dimensions changed" point (sift.c:651-656) -- and replayed from then on.
    bool use_graph = true;              // $SIFT3D_CUDA_NO_GRAPH switches it off
    cudaGraphExec_t graph_exec = nullptr;
    double graph_key[13] = {};
    bool counted = false;               // in g_live_ctx
    int graph_key_runs = 0;             // eager runs seen with graph_key
    uint64_t graph_launches = 0;        // kernel launches inside the graph
    uint64_t plan_gen = 0;              // bumped whenever device buffers may have moved
    DevBuf orient_plan;                 // OrientPlan on the device
    DevBuf orient_tab, orient_ntab;     // orientation window tables (OrientEntry) and their sizes
    double orient_key[11] = {};         // geometry the tables were built for
    bool orient_tab_ok = false;
    const OrientEntry *orient_tab_ptr[SIFT3D_CUDA_MAX_OCTAVES][SIFT3D_CUDA_MAX_FILTERS] = {};
    OrientPlan *h_plan = nullptr;       // pinned staging copy
    OrientPlan h_plan_dev;              // what the device copy holds (valid if plan_uploaded)
    bool plan_uploaded = false;
    uint64_t plan_uploaded_gen = 0;
    DevBuf work;                        // work counter of the persistent orientation kernel
    DevBuf taps;                        // (1 + filters) * SIFT3D_CUDA_MAX_TAPS floats
    int base_octave = 0;                // octave number of this context's octave 0
    DevBuf scalars;                     // [0] image max-abs; [1 + o*MAX_FILTERS + s] DoG max
    DevBuf cand_raw, cand_sorted;       // Cand regions, one per octave
    DevBuf cand_grouped, order_ints;    // scratch of the bucket ordering of busy regions
    OrderScratch order_scratch;
    DevBuf counts;                      // int per octave
    DevBuf kps;                         // sift3d_cuda_keypoint regions, one per octave (device)
    // ... or, when they fit kKpsHostMax, in page-locked host memory that the orientation
    // kernel writes directly (88 B per candidate over PCIe, under the kernel): the host reads
    // the records in place after the one synchronisation of detect -- no device-to-host copy
    // and no second synchronisation (48 us + ~30 us per 256^3 volume).
    sift3d_cuda_keypoint *kps_host = nullptr;
    size_t kps_host_bytes = 0;
    sift3d_cuda_candidate_flag *kflags_host = nullptr;  // behind the records in the same allocation
    sift3d_cuda_keypoint *kps_dev() const {
        return kps_host ? kps_host : (sift3d_cuda_keypoint *) kps.p;  // UVA: same address on the device
    }
    std::vector<int> cand_base, cand_cap, cand_min;
    int *h_counts = nullptr;            // pinned
    bool have_pyramid = false;
    int total_candidates = 0;

    // describe
    DevBuf desc_jobs, desc_out;
    DescJob *h_jobs = nullptr;          // pinned (device-readable) staging of the job list
    size_t h_jobs_cap = 0;
    sift3d_cuda_timings tm;

    // user marks + per-blur profiling
    cudaEvent_t marks[8] = {};
    cudaEvent_t sleep_ev = nullptr;     // $SIFT3D_CUDA_BLOCKING_SYNC: the two host waits of a step sleep
    bool profiling = false;
    std::vector<cudaEvent_t> blur_ev;    // 2 per level: [2 * (o * nlev + s + 1)] start, +1 stop
    std::vector<int> blur_launches;      // launches per blur in the last detect

    Level &level(int o, int s) { return levels[(size_t) o * nlev + (s + 1)]; }
    float *d_taps(int filter /* -1 = first */) {
        return (float *) taps.p + (size_t) (filter + 1) * SIFT3D_CUDA_MAX_TAPS;
    }
    float *d_scalar(int i) { return (float *) scalars.p + i; }
};

static int set_device(sift3d_cuda_ctx *ctx) {
    S3D_CK(cudaSetDevice(ctx->device));
    return 0;
}

extern "C" {

int sift3d_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

// Contexts alive per device (pipelines sharing a GPU, see detect_impl).
static std::atomic<int> g_live_ctx[SIFT3D_CUDA_MULTI_MAX_DEVICES];
static int live_contexts(int device) {
    return device >= 0 && device < SIFT3D_CUDA_MULTI_MAX_DEVICES ? g_live_ctx[device].load() : 1;
}

sift3d_cuda_ctx *sift3d_cuda_ctx_create(int device) {
    int n = sift3d_cuda_device_count();
    if (n <= 0) {
        S3D_ERR("sift3d_cuda: no CUDA device available -- this library has no CPU fallback\n");
        return nullptr;
    }
    if (device < 0) {
        const char *e = getenv("SIFT3D_CUDA_DEVICE");
        device = e ? atoi(e) : 0;
    }
    if (device >= n) {
        S3D_ERR("sift3d_cuda: device %d requested but only %d visible\n", device, n);
        return nullptr;
    }
    sift3d_cuda_ctx *ctx = new (std::nothrow) sift3d_cuda_ctx();
    if (!ctx) return nullptr;
    ctx->device = device;
    memset(&ctx->tm, 0, sizeof(ctx->tm));
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        S3D_ERR("sift3d_cuda: cannot initialise device %d: %s\n", device,
                cudaGetErrorString(cudaGetLastError()));
        delete ctx;
        return nullptr;
    }
    for (int i = 0; i < EV_COUNT; i++) cudaEventCreate(&ctx->ev[i]);
    ctx->serial = getenv("SIFT3D_CUDA_SERIAL") != nullptr;
    ctx->use_graph = getenv("SIFT3D_CUDA_NO_GRAPH") == nullptr;
    {
        const char *e = getenv("SIFT3D_CUDA_EXACT");
        ctx->exact = e != nullptr && atoi(e) != 0;
    }
    for (int o = 0; o < SIFT3D_CUDA_MAX_OCTAVES; o++) {
        cudaEventCreateWithFlags(&ctx->ev_seed[o], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ctx->ev_done[o], cudaEventDisableTiming);
    }
    for (int i = 0; i < 8; i++) cudaEventCreate(&ctx->marks[i]);
    for (int i = 0; i < DV_COUNT; i++) cudaEventCreate(&ctx->dv[i]);
    if (cudaMallocHost((void **) &ctx->h_counts, sizeof(int) * 64) != cudaSuccess ||
        cudaMallocHost((void **) &ctx->h_plan, sizeof(OrientPlan)) != cudaSuccess ||
        upload_mesh() != 0) {
        S3D_ERR("sift3d_cuda: context setup failed\n");
        sift3d_cuda_ctx_destroy(ctx);
        return nullptr;
    }
    if (ctx->device >= 0 && ctx->device < SIFT3D_CUDA_MULTI_MAX_DEVICES) {
        g_live_ctx[ctx->device]++;
        ctx->counted = true;
    }
    return ctx;
}

void sift3d_cuda_ctx_destroy(sift3d_cuda_ctx *ctx) {
    if (!ctx) return;
    if (ctx->counted) g_live_ctx[ctx->device]--;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->graph_exec) cudaGraphExecDestroy(ctx->graph_exec);
    ctx->graph_exec = nullptr;
    ctx->pyr.release();
    ctx->input.release();
    ctx->tmp_a.release();
    ctx->tmp_b.release();
    ctx->tmp_side.release();
    ctx->orient_plan.release();
    ctx->orient_tab.release();
    ctx->orient_ntab.release();
    ctx->work.release();
    for (int o = 0; o < SIFT3D_CUDA_MAX_OCTAVES; o++) {
        if (ctx->oct_stream[o]) {
            cudaStreamSynchronize(ctx->oct_stream[o]);
            cudaStreamDestroy(ctx->oct_stream[o]);
        }
        if (ctx->ev_seed[o]) cudaEventDestroy(ctx->ev_seed[o]);
        if (ctx->ev_done[o]) cudaEventDestroy(ctx->ev_done[o]);
    }
    if (ctx->h_plan) cudaFreeHost(ctx->h_plan);
    if (ctx->h_jobs) cudaFreeHost(ctx->h_jobs);
    ctx->taps.release();
    ctx->scalars.release();
    ctx->cand_raw.release();
    ctx->cand_sorted.release();
    ctx->cand_grouped.release();
    ctx->order_ints.release();
    ctx->counts.release();
    ctx->kps.release();
    if (ctx->kps_host) cudaFreeHost(ctx->kps_host);
    ctx->kps_host = nullptr;
    ctx->desc_jobs.release();
    ctx->desc_out.release();
    for (int i = 0; i < 8; i++)
        if (ctx->marks[i]) cudaEventDestroy(ctx->marks[i]);
    if (ctx->sleep_ev) cudaEventDestroy(ctx->sleep_ev);
    for (cudaEvent_t e : ctx->blur_ev) cudaEventDestroy(e);
    if (ctx->h_counts) cudaFreeHost(ctx->h_counts);
    for (int i = 0; i < EV_COUNT; i++)
        if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
    for (int i = 0; i < DV_COUNT; i++)
        if (ctx->dv[i]) cudaEventDestroy(ctx->dv[i]);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

void *sift3d_cuda_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (bytes == 0) bytes = 1;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}

void sift3d_cuda_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

// Candidate regions, one per octave: room for 1/32 of the octave's voxels (typical images
// yield 0.04 - 0.4 %), or for what an earlier detect on this plan was seen to need.
static void detect_timings(sift3d_cuda_ctx *ctx);

// The host waits at the end of detect and of describe. By default the CUDA runtime spins, which is
// the lowest latency as long as every waiting thread has a core; with more pipelines than cores
// in the box (8 ranks x 4 pipelines on 16 hardware threads) the spinning threads take the cores
// from the ones that have launches and uploads to issue: $SIFT3D_CUDA_BLOCKING_SYNC=1 makes the
// waits sleep on an event instead.
static int host_wait(sift3d_cuda_ctx *ctx, cudaStream_t st) {
    static const bool sleepy = getenv("SIFT3D_CUDA_BLOCKING_SYNC") != nullptr &&
                               atoi(getenv("SIFT3D_CUDA_BLOCKING_SYNC")) != 0;
    if (!sleepy) {
        S3D_CK(cudaStreamSynchronize(st));
        return 0;
    }
    if (!ctx->sleep_ev)
        S3D_CK(cudaEventCreateWithFlags(&ctx->sleep_ev, cudaEventBlockingSync | cudaEventDisableTiming));
    S3D_CK(cudaEventRecord(ctx->sleep_ev, st));
    S3D_CK(cudaEventSynchronize(ctx->sleep_ev));
    return 0;
}

static int plan_candidates(sift3d_cuda_ctx *ctx) {
    const int noct = ctx->desc.num_octaves;
    ctx->plan_gen++;  // buffers may move: a captured graph holds their addresses
    ctx->cand_base.assign(noct, 0);
    ctx->cand_cap.assign(noct, 0);
    size_t cand_total = 0;
    for (int o = 0; o < noct; o++) {
        const size_t nv = ctx->level(o, 0).nvox();
        size_t cap = nv / 32;
        if (cap < 4096) cap = 4096;
        if (cap < (size_t) ctx->cand_min[o]) cap = (size_t) ctx->cand_min[o];
        if (cap > nv) cap = nv;
        if (cap > ((size_t) 1 << 27)) cap = (size_t) 1 << 27;
        ctx->cand_base[o] = (int) cand_total;
        ctx->cand_cap[o] = (int) cap;
        cand_total += cap;
    }
    {
        // bucket ordering (extrema.cu): one bucket per (keypoint level, plane) of an octave
        OrderScratch &S = ctx->order_scratch;
        memset(&S, 0, sizeof(S));
        size_t ints = 0;
        for (int o = 0; o < noct; o++) {
            S.nz[o] = ctx->level(o, 0).nz;
            S.nbkt[o] = ctx->desc.num_kp_levels * S.nz[o];
            S.off[o] = ints;
            ints += 3 * (size_t) S.nbkt[o];
        }
        if (ctx->cand_grouped.ensure(cand_total * sizeof(Cand)) || ctx->order_ints.ensure((ints + 1) * sizeof(int)))
            return -1;
        S.grouped = (Cand *) ctx->cand_grouped.p;
        S.ints = (int *) ctx->order_ints.p;
    }
    if (ctx->cand_raw.ensure(cand_total * sizeof(Cand)) ||
        ctx->cand_sorted.ensure(cand_total * sizeof(Cand)) ||
        ctx->counts.ensure(sizeof(int) * 64))
        return -1;
    static const bool no_host_kps = getenv("SIFT3D_CUDA_NO_HOST_KPS") != nullptr;
    const size_t rbytes = cand_total * sizeof(sift3d_cuda_keypoint);  // multiple of 8
    const size_t kbytes = rbytes + cand_total * sizeof(sift3d_cuda_candidate_flag);
    constexpr size_t kKpsHostMax = (size_t) 256 << 20;
    if (!no_host_kps && kbytes <= kKpsHostMax) {
        if (kbytes > ctx->kps_host_bytes) {
            if (ctx->kps_host) cudaFreeHost(ctx->kps_host);
            ctx->kps_host = nullptr;
            ctx->kps_host_bytes = 0;
            void *hp = nullptr;
            if (cudaHostAlloc(&hp, kbytes, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
                void *dp = nullptr;
                if (cudaHostGetDevicePointer(&dp, hp, 0) == cudaSuccess && dp == hp) {
                    ctx->kps_host = (sift3d_cuda_keypoint *) hp;
                    ctx->kps_host_bytes = kbytes;
                } else {
                    cudaFreeHost(hp);
                }
            }
            cudaGetLastError();
        }
        if (ctx->kps_host) {
            ctx->kflags_host = (sift3d_cuda_candidate_flag *) ((char *) ctx->kps_host + rbytes);
            ctx->kps.release();
            return 0;
        }
    } else if (ctx->kps_host) {
        cudaFreeHost(ctx->kps_host);
        ctx->kps_host = nullptr;
        ctx->kps_host_bytes = 0;
    }
    ctx->kflags_host = nullptr;
    if (ctx->kps.ensure(rbytes)) return -1;
    return 0;
}

int sift3d_cuda_plan(sift3d_cuda_ctx *ctx, const sift3d_cuda_plan_desc *d) {
    if (!ctx || !d) return -1;
    if (set_device(ctx)) return -1;
    if (d->num_octaves < 1 || d->num_octaves > SIFT3D_CUDA_MAX_OCTAVES || d->num_kp_levels < 1 ||
        d->num_kp_levels + 2 > SIFT3D_CUDA_MAX_FILTERS || d->nx < 1 || d->ny < 1 || d->nz < 1) {
        S3D_ERR("sift3d_cuda_plan: invalid geometry\n");
        return -1;
    }
    if (d->nx >= (1 << 20) || d->ny >= (1 << 20) || d->nz >= (1 << 20)) {
        S3D_ERR("sift3d_cuda_plan: dimensions must be below 2^20\n");
        return -1;
    }
    S3D_CK(cudaStreamSynchronize(ctx->stream));
    ctx->planned = false;
    ctx->have_pyramid = false;
    ctx->orient_tab_ok = false;  // the window tables depend on the plan's units and scales
    ctx->desc = *d;
    ctx->nlev = d->num_kp_levels + 3;
    const int nlev = ctx->nlev, noct = d->num_octaves;

    // level geometry: dims halve per octave (imutil.c:1545-1548)
    ctx->levels.assign((size_t) noct * nlev, Level());
    for (cudaEvent_t e : ctx->blur_ev) cudaEventDestroy(e);
    ctx->blur_ev.assign((size_t) 2 * noct * nlev, nullptr);
    for (cudaEvent_t &e : ctx->blur_ev) S3D_CK(cudaEventCreate(&e));
    ctx->blur_launches.assign((size_t) noct * nlev, 0);
    size_t total = 0;
    {
        int nx = d->nx, ny = d->ny, nz = d->nz;
        for (int o = 0; o < noct; o++) {
            if (nx < 1 || ny < 1 || nz < 1) {
                S3D_ERR("sift3d_cuda_plan: too many octaves for these dimensions\n");
                return -1;
            }
            for (int s = -1; s < nlev - 1; s++) {
                Level &L = ctx->level(o, s);
                L.nx = nx; L.ny = ny; L.nz = nz;
                L.scale = d->scales[o][s + 1];
                total += (L.nvox() + 63) & ~(size_t) 63;  // 256-byte aligned levels
            }
            nx /= 2; ny /= 2; nz /= 2;
        }
    }
    if (ctx->pyr.ensure(total * sizeof(float))) return -1;
    {
        size_t off = 0;
        for (int o = 0; o < noct; o++)
            for (int s = -1; s < nlev - 1; s++) {
                Level &L = ctx->level(o, s);
                L.d = (float *) ctx->pyr.p + off;
                off += (L.nvox() + 63) & ~(size_t) 63;
            }
    }
    const size_t n0 = (size_t) d->nx * d->ny * d->nz;
    if (ctx->input.ensure(n0 * sizeof(float)) || ctx->tmp_a.ensure(n0 * sizeof(float)) ||
        ctx->tmp_b.ensure(n0 * sizeof(float)))
        return -1;

    // octaves >= 1: own stream and own 3-pass scratch each
    {
        ctx->side_off.assign(noct, 0);
        size_t side_total = 0;
        for (int o = 1; o < noct; o++) {
            ctx->side_off[o] = side_total;
            side_total += 2 * ((ctx->level(o, 0).nvox() + 63) & ~(size_t) 63);
            // higher priority than the octave-0 stream: the small octaves are the tail of the
            // dependency chain, so their kernels go first whenever SMs free up
            int prio_lo = 0, prio_hi = 0;
            cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
            // (numerically lower = higher priority) octave 1 above octave 0, the rest above both
            const int prio = o == 1 ? (prio_hi + prio_lo) / 2 : prio_hi;
            if (!ctx->oct_stream[o] &&
                cudaStreamCreateWithPriority(&ctx->oct_stream[o], cudaStreamNonBlocking, prio) !=
                    cudaSuccess) {
                S3D_ERR("sift3d_cuda_plan: cannot create a stream\n");
                return -1;
            }
        }
        if (ctx->tmp_side.ensure((side_total + 64) * sizeof(float)) ||
            ctx->orient_plan.ensure(sizeof(OrientPlan)) || ctx->work.ensure(sizeof(int) * 64))
            return -1;
    }

    // filters
    const int nfilt = d->num_kp_levels + 2;
    if (ctx->taps.ensure((size_t) (nfilt + 1) * SIFT3D_CUDA_MAX_TAPS * sizeof(float))) return -1;
    if (d->first.width < 1 || d->first.width > SIFT3D_CUDA_MAX_TAPS) {
        S3D_ERR("sift3d_cuda_plan: unsupported filter width %d\n", d->first.width);
        return -1;
    }
    S3D_CK(cudaMemcpyAsync(ctx->d_taps(-1), d->first.taps, sizeof(float) * d->first.width,
                           cudaMemcpyHostToDevice, ctx->stream));
    for (int i = 0; i < nfilt; i++) {
        if (d->octave[i].width < 1 || d->octave[i].width > SIFT3D_CUDA_MAX_TAPS) {
            S3D_ERR("sift3d_cuda_plan: unsupported filter width %d\n", d->octave[i].width);
            return -1;
        }
        S3D_CK(cudaMemcpyAsync(ctx->d_taps(i), d->octave[i].taps,
                               sizeof(float) * d->octave[i].width, cudaMemcpyHostToDevice,
                               ctx->stream));
    }
    if (ctx->scalars.ensure(sizeof(float) * (1 + (size_t) noct * SIFT3D_CUDA_MAX_FILTERS)))
        return -1;

    // candidate regions
    ctx->cand_min.assign(noct, 0);
    if (plan_candidates(ctx)) return -1;
    S3D_CK(cudaStreamSynchronize(ctx->stream));
    ctx->planned = true;
    return 0;
}

// The whole detection pipeline. Three ways in:
//   * plain: image -> all octaves (sift3d_cuda_detect);
//   * seed:  `image` is level -1 of this context's octave 0, already blurred and decimated
//            by a bigger pyramid (sift3d_cuda_detect_from_seed): no im_scale, no first blur;
//   * slab:  this context holds a z-slab (+ halo planes) of a taller volume
//            (sift3d_cuda_detect_slab): octaves 0 .. slab->num_sharded_octaves-1 produce only
//            the planes this rank owns, the halo planes of every level are filled by the
//            caller's exchange callback and the global max reductions by its reduce callback.
static int detect_impl(sift3d_cuda_ctx *ctx, const float *image, int image_on_device,
                       const double units[3], double peak_thresh, double corner_thresh,
                       int *num_candidates, bool from_seed, const sift3d_cuda_slab_desc *slab,
                       const sift3d_cuda_slab_callbacks *cb, int attempt = 0) {
    if (!ctx || !ctx->planned || !image || !units) {
        S3D_ERR("sift3d_cuda_detect: context not planned or NULL argument\n");
        return -1;
    }
    if (set_device(ctx)) return -1;
    cudaStream_t st = ctx->stream;
    const sift3d_cuda_plan_desc &D = ctx->desc;
    const int nlev = ctx->nlev, kp = D.num_kp_levels;
    const int noct_plan = D.num_octaves;
    // octaves processed by this call
    const int noct = slab ? slab->num_sharded_octaves : noct_plan;
    const size_t n0 = (size_t) D.nx * D.ny * D.nz;
    ctx->have_pyramid = false;
    if (slab) {
        if (!cb || !cb->reduce_max || !cb->exchange || noct < 1 || noct > noct_plan) {
            S3D_ERR("sift3d_cuda_detect_slab: bad slab description\n");
            return -1;
        }
        for (int o = 0; o < noct_plan && o <= noct; o++) {
            const Level &L = ctx->level(o, 0);
            if (slab->own0[o] < 0 || slab->own1[o] > L.nz || slab->own0[o] >= slab->own1[o] ||
                slab->org[o] < 0) {
                S3D_ERR("sift3d_cuda_detect_slab: bad owned range of octave %d\n", o);
                return -1;
            }
        }
    }
    const bool serial = ctx->serial || slab != nullptr;
    // planes [zr0(o), zr1(o)) of octave o are produced / scanned
    auto zr0 = [&](int o) { return slab ? slab->own0[o] : 0; };
    auto zr1 = [&](int o) { return slab ? slab->own1[o] : ctx->level(o, 0).nz; };
    auto zorg = [&](int o) { return slab ? slab->org[o] : 0; };
    // hand the stream to the caller's collective (host callbacks, so the stream is drained)
    auto host_sync = [&]() -> int {
        if (cb != nullptr && cb->in_stream) return 0;  // the callbacks enqueue behind the stream
        S3D_CK(cudaStreamSynchronize(st));
        return 0;
    };

    // Units: octave 0 takes the units of this image (apply_Sep_FIR_filter copies them from
    // the source, imutil.c:1145); octaves >= 1 keep what resize_Pyramid assigned when the
    // dims last changed, doubled per octave (imutil.c:1545-1548, sift.c:651-656).
    for (int o = 0; o < noct_plan; o++)
        for (int s = -1; s < nlev - 1; s++) {
            Level &L = ctx->level(o, s);
            const double *u = o == 0 ? units : D.units;
            L.ux = ldexp(u[0], o);
            L.uy = ldexp(u[1], o);
            L.uz = ldexp(u[2], o);
        }

    S3D_CK(cudaEventRecord(ctx->ev[EV_START], st));
    const float *d_in = image;
    if (from_seed) {
        // level -1 of octave 0 is given (sift.c:694-705 produced it in the bigger pyramid)
        S3D_CK(cudaMemcpyAsync(ctx->level(0, -1).d, image, n0 * sizeof(float),
                               image_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    } else if (!image_on_device) {
        size_t first = 0, count = n0;
        if (slab && cb->in_stream && cb->input_reach > 0) {
            // only the planes the first blur reads around the owned range
            const size_t plane = (size_t) D.nx * D.ny;
            const int p0 = slab->own0[0] - cb->input_reach > 0 ? slab->own0[0] - cb->input_reach : 0;
            const int p1 = slab->own1[0] + cb->input_reach < D.nz ? slab->own1[0] + cb->input_reach : D.nz;
            first = (size_t) p0 * plane;
            count = (size_t) (p1 - p0) * plane;
        }
        S3D_CK(cudaMemcpyAsync((float *) ctx->input.p + first, image + first, count * sizeof(float),
                               cudaMemcpyHostToDevice, st));
        d_in = (const float *) ctx->input.p;
    }
    S3D_CK(cudaEventRecord(ctx->ev[EV_H2D], st));

    // im_scale: max |v| (imutil.c:681-695); the division is applied on load by the first blur.
    if (slab) S3D_CK(cudaMemsetAsync(ctx->scalars.p, 0, ctx->scalars.bytes, st));
    if (!from_seed) {
        const size_t plane = (size_t) D.nx * D.ny;
        if (launch_absmax(st, d_in + (size_t) zr0(0) * plane, (size_t) (zr1(0) - zr0(0)) * plane,
                          ctx->d_scalar(0), &ctx->launches))
            return -1;
        if (slab) {  // the maximum is over the whole volume
            if (host_sync()) return -1;
            if (cb->reduce_max(cb->user, ctx->d_scalar(0), 1)) return -1;
        }
    }
    S3D_CK(cudaMemsetAsync(ctx->counts.p, 0, sizeof(int) * 64, st));
    S3D_CK(cudaMemsetAsync(ctx->work.p, 0, sizeof(int) * 64, st));
    // what the per-candidate kernels need to know about each octave (read by every octave's
    // orientation pass, so it is uploaded before any side stream starts)
    {
        OrientPlan &H = *ctx->h_plan;
        memset(&H, 0, sizeof(H));
        H.noct = noct;
        for (int o = 0; o < noct; o++) {
            OctaveInfo &info = H.oct[o];
            for (int s = 0; s < kp; s++) {
                info.lvl[s] = ctx->level(o, s).d;
                info.sd[s] = ctx->level(o, s).scale;
            }
            Level &L = ctx->level(o, 0);
            info.o = o + ctx->base_octave;
            info.nx = L.nx; info.ny = L.ny; info.nz = L.nz;
            info.ux = (float) L.ux; info.uy = (float) L.uy; info.uz = (float) L.uz;
            H.base[o] = ctx->cand_base[o];
            H.cap[o] = ctx->cand_cap[o];
        }
        // Orientation window tables: rebuilt only when the geometry they depend on changes
        // (units and level scales; one tiny launch), otherwise the cached pointers are reused.
        static const bool no_tab = getenv("SIFT3D_CUDA_NO_ORIENT_TABLES") != nullptr;
        if (!no_tab) {
            const Level &L0 = ctx->level(0, 0);
            // octave 0 takes the image's units, octaves >= 1 the plan's (see above): both are
            // part of the key, or a detect with new units on unchanged dims followed by a
            // re-plan would reuse tables built for the old octave >= 1 spacing
            const double key[11] = {(double) noct, (double) kp, L0.ux, L0.uy, L0.uz, L0.scale,
                                    ctx->level(noct - 1, kp - 1).scale, (double) ctx->base_octave,
                                    D.units[0], D.units[1], D.units[2]};
            if (!ctx->orient_tab_ok || memcmp(key, ctx->orient_key, sizeof(key)) != 0) {
                ctx->orient_tab_ok = false;
                size_t off[SIFT3D_CUDA_MAX_OCTAVES][SIFT3D_CUDA_MAX_FILTERS];
                const size_t cap = orient_table_capacity(H, kp, off);
                if (cap > 0 && cap * sizeof(OrientEntry) <= ((size_t) 256 << 20)) {
                    if (ctx->orient_tab.ensure(cap * sizeof(OrientEntry)) ||
                        ctx->orient_ntab.ensure(sizeof(int) * SIFT3D_CUDA_MAX_OCTAVES * SIFT3D_CUDA_MAX_FILTERS))
                        return -1;
                    const int r = launch_orient_tables(st, H, kp, (OrientEntry *) ctx->orient_tab.p,
                                                       (int *) ctx->orient_ntab.p, &ctx->launches);
                    if (r < 0) return -1;
                    if (r == 0) {
                        memcpy(ctx->orient_tab_ptr, H.tab, sizeof(H.tab));
                        memcpy(ctx->orient_key, key, sizeof(key));
                        ctx->orient_tab_ok = true;
                    }
                }
            }
            if (ctx->orient_tab_ok) {
                memcpy(H.tab, ctx->orient_tab_ptr, sizeof(H.tab));
                H.ntab = (const int *) ctx->orient_ntab.p;
            } else {
                memset(H.tab, 0, sizeof(H.tab));
                H.ntab = nullptr;
            }
        }
        // The plan changes only with the geometry, the units or the tables: it is uploaded
        // when it differs from the device copy. (Every detect used to upload it; a host-to-device
        // copy queues on the copy engine BEHIND an image upload in flight for the next volume --
        // sift3d_cuda_upload_begin -- and held the whole detect back for the length of that
        // upload, 1.2 ms for 64 MiB.)
        if (!ctx->plan_uploaded || ctx->plan_uploaded_gen != ctx->plan_gen ||
            memcmp(&H, &ctx->h_plan_dev, sizeof(H)) != 0) {
            S3D_CK(cudaMemcpyAsync(ctx->orient_plan.p, &H, sizeof(H), cudaMemcpyHostToDevice, st));
            // (H is the pinned staging copy: it must not change before the copy has run)
            S3D_CK(cudaStreamSynchronize(st));
            memcpy(&ctx->h_plan_dev, &H, sizeof(H));
            ctx->plan_uploaded = true;
            ctx->plan_uploaded_gen = ctx->plan_gen;
        }
    }
    S3D_CK(cudaEventRecord(ctx->ev[EV_SCALE], st));

    // build_gpyr (sift.c:662-711) + detect_extrema (sift.c:735-871), one octave per stream:
    // octave o + 1 starts as soon as its seed level (2x decimation of level ds_level of octave
    // o, fused into that blur's epilogue) exists, while octave o finishes its upper levels.
    const int ds_level = (nlev - 2) - 2 > -1 ? (nlev - 2) - 2 : -1;  // sift.c:696-697
    // Octaves of at most kTinyMaxVox voxels (and everything after them) run as one chain
    // kernel on the stream of the first of them.
    int tiny_from = -1;
    static const bool no_tiny = getenv("SIFT3D_CUDA_NO_TINY") != nullptr;
    static const int tiny_vox_env = getenv("SIFT3D_CUDA_TINY_VOX") ? atoi(getenv("SIFT3D_CUDA_TINY_VOX")) : kTinyMaxVox;
    const int tiny_vox = tiny_vox_env < kTinyMaxVox ? tiny_vox_env : kTinyMaxVox;
    if (!slab && !no_tiny) {
        for (int o = from_seed ? 0 : 1; o < noct; o++)
            if (ctx->level(o, 0).nvox() <= (size_t) tiny_vox) {
                if (noct - o <= kTinyMaxOct) tiny_from = o;
                break;
            }
    }
    auto stream_of = [&](int o) {
        if (o == 0 || serial) return st;
        return ctx->oct_stream[tiny_from >= 0 && o > tiny_from ? tiny_from : o];
    };
    auto run_tiny = [&](cudaStream_t so) -> int {
        TinyChainParams T;
        memset(&T, 0, sizeof(T));
        T.noct = noct - tiny_from;
        T.nfilt = nlev - 1;
        T.ds_level = ds_level;
        static const bool no_band = getenv("SIFT3D_CUDA_NO_BAND") != nullptr;
        T.banded = !ctx->exact && !no_band && T.noct <= kTinyMaxOct ? 1 : 0;
        for (int i = 0; i < T.nfilt; i++) {
            T.taps[i] = ctx->d_taps(i);
            T.hw[i] = D.octave[i].width / 2;
        }
        for (int t = 0; t < T.noct; t++) {
            const int o = tiny_from + t;
            const Level &L = ctx->level(o, 0);
            TinyOctave &O = T.oct[t];
            O.nx = L.nx; O.ny = L.ny; O.nz = L.nz;
            tiny_axis(L.nx, L.ux, &O.uf[0], &O.sh[0]);
            tiny_axis(L.ny, L.uy, &O.uf[1], &O.sh[1]);
            tiny_axis(L.nz, L.uz, &O.uf[2], &O.sh[2]);
            for (int i = 0; i <= T.nfilt; i++) O.lvl[i] = ctx->level(o, i - 1).d;
            if (T.banded) {
                const double un[3] = {L.ux, L.uy, L.uz};
                const int dn[3] = {L.nx, L.ny, L.nz};
                for (int i = 0; i < T.nfilt && T.banded; i++)
                    for (int a = 0; a < 3 && T.banded; a++) {
                        const int r = tiny_band(dn[a], un[a], D.octave[i], &T.band[t][i][a]);
                        if (r < 0) return -1;
                        if (r > 0) T.banded = 0;  // not representable: the literal chain
                    }
            }
            for (int i = 0; i < T.nfilt; i++)
                O.dogmax[i] = (i - 1 >= 0 && i - 1 < kp)
                                  ? ctx->d_scalar(1 + o * SIFT3D_CUDA_MAX_FILTERS + (i - 1)) : nullptr;
        }
        return launch_tiny_octaves(so, T, &ctx->launches);
    };
    auto run_level = [&](int o, int s) -> int {  // blur that produces level (o, s)
        cudaStream_t so = stream_of(o);
        if (tiny_from >= 0 && o >= tiny_from) {
            const size_t li = (size_t) o * nlev + (s + 1);
            if (ctx->profiling) S3D_CK(cudaEventRecord(ctx->blur_ev[2 * li], so));
            const bool first = o == tiny_from && s == 0;
            if (first && run_tiny(so)) return -1;
            if (ctx->profiling) S3D_CK(cudaEventRecord(ctx->blur_ev[2 * li + 1], so));
            ctx->blur_launches[li] = first ? 1 : 0;
            return 0;
        }
        float *ta, *tb;
        if (o == 0) {
            ta = (float *) ctx->tmp_a.p;
            tb = (float *) ctx->tmp_b.p;
        } else {
            ta = (float *) ctx->tmp_side.p + ctx->side_off[o];
            tb = ta + ((ctx->level(o, 0).nvox() + 63) & ~(size_t) 63);
        }
        Level &L = ctx->level(o, s);
        const size_t li = (size_t) o * nlev + (s + 1);
        const uint64_t l0 = ctx->launches;
        if (ctx->profiling) S3D_CK(cudaEventRecord(ctx->blur_ev[2 * li], so));
        // the level that seeds the next octave is decimated as it is produced
        float *ds_dst = (o != noct_plan - 1 && s == ds_level) ? ctx->level(o + 1, -1).d : nullptr;
        int rc;
        if (s == -1) {
            const double u[3] = {L.ux, L.uy, L.uz};
            rc = launch_blur(so, d_in, L.d, ta, tb, L.nx, L.ny, L.nz, u, D.first, ctx->d_taps(-1),
                             ctx->d_scalar(0), nullptr, nullptr, ds_dst, false, &ctx->launches,
                             zr0(o), zr1(o), zorg(o), ctx->exact, slab ? slab->vol_nz[o] : 0);
        } else {
            Level &P = ctx->level(o, s - 1);
            const double u[3] = {P.ux, P.uy, P.uz};
            // DoG level s-1 = G(s-1) - G(s); its max-abs is needed for keypoint levels only
            float *slot = nullptr;
            if (s - 1 >= 0 && s - 1 < kp) slot = ctx->d_scalar(1 + o * SIFT3D_CUDA_MAX_FILTERS + (s - 1));
            rc = launch_blur(so, P.d, L.d, ta, tb, L.nx, L.ny, L.nz, u, D.octave[s], ctx->d_taps(s),
                             nullptr, slot, P.d, ds_dst, false, &ctx->launches, zr0(o), zr1(o),
                             zorg(o), ctx->exact, slab ? slab->vol_nz[o] : 0);
        }
        if (rc) return -1;
        if (ctx->profiling) S3D_CK(cudaEventRecord(ctx->blur_ev[2 * li + 1], so));
        ctx->blur_launches[li] = (int) (ctx->launches - l0);
        if (ds_dst != nullptr && !serial) S3D_CK(cudaEventRecord(ctx->ev_seed[o], so));
        if (slab) {
            // halo planes of the new level (and of the seed it decimated into) come from the
            // neighbouring ranks, which own them
            if (host_sync()) return -1;
            if (cb->exchange(cb->user, o, s, L.d, L.nx, L.ny, L.nz, zr0(o), zr1(o))) return -1;
            if (ds_dst != nullptr) {
                const Level &N = ctx->level(o + 1, -1);
                if (cb->exchange(cb->user, o + 1, -1, N.d, N.nx, N.ny, N.nz, zr0(o + 1), zr1(o + 1)))
                    return -1;
            }
        }
        return 0;
    };
    // Extrema of octave o, keypoint levels [s_begin, s_end)
    auto run_extrema = [&](int o, int s_begin, int s_end) -> int {
        const float *g[SIFT3D_CUDA_MAX_FILTERS + 1];
        for (int s = -1; s < nlev - 1; s++) g[s + 1] = ctx->level(o, s).d;
        Level &L = ctx->level(o, 0);
        return launch_extrema(stream_of(o), g, kp, L.nx, L.ny, L.nz,
                              ctx->d_scalar(1 + o * SIFT3D_CUDA_MAX_FILTERS), peak_thresh,
                              (Cand *) ctx->cand_raw.p + ctx->cand_base[o], ctx->cand_cap[o],
                              (int *) ctx->counts.p + o, &ctx->launches, zr0(o), zr1(o), s_begin, s_end);
    };
    // Keypoint levels of octave 0 whose four Gaussian levels exist once the seed level does
    // (level s needs G(s-1) .. G(s+2)): scanned while octave 0 waits for octave 1's seed
    // chain, instead of on the critical path after the last blur.
    const int early_levels = (!slab && noct > 1) ? (ds_level - 1 > 0 ? (ds_level - 1 < kp ? ds_level - 1 : kp) : 0) : 0;
    // ---- CUDA graph of everything behind the first blur (see the context's graph fields)
    static const int reserve_env = getenv("SIFT3D_CUDA_RESERVE") ? atoi(getenv("SIFT3D_CUDA_RESERVE")) : 14;
    static const int reserve_first = getenv("SIFT3D_CUDA_RESERVE_FIRST") ? atoi(getenv("SIFT3D_CUDA_RESERVE_FIRST")) : 40;
    const bool graph_eligible = ctx->use_graph && !slab && !from_seed && !serial && !ctx->profiling && attempt == 0;
    // other detectors (pipelines) live on this device: latency-bound persistent kernels then keep
    // their grids small so that the other pipelines' kernels fit next to them (launch_orient_all)
    const bool shared_device = live_contexts(ctx->device) > 1;
    bool replay = false, capturing = false;
    if (graph_eligible) {
        const double key[13] = {(double) ctx->plan_gen, units[0], units[1], units[2], peak_thresh, corner_thresh,
                                ctx->exact ? 1.0 : 0.0, (double) ctx->base_octave, ctx->orient_tab_ok ? 1.0 : 0.0,
                                (double) (uintptr_t) ctx->orient_tab.p, (double) reserve_env, (double) reserve_first,
                                shared_device ? 1.0 : 0.0};
        if (memcmp(key, ctx->graph_key, sizeof(key)) != 0) {
            if (ctx->graph_exec) cudaGraphExecDestroy(ctx->graph_exec);
            ctx->graph_exec = nullptr;
            memcpy(ctx->graph_key, key, sizeof(key));
            ctx->graph_key_runs = 0;
        }
        replay = ctx->graph_exec != nullptr;
        // the first run with a key is eager (one-time kernel attribute set-up happens there)
        capturing = !replay && ctx->graph_key_runs >= 1;
        ctx->graph_key_runs++;
    }
    // timing marks: inside a capture they become event-record nodes of the graph
    auto mark = [&](cudaEvent_t ev) -> int {
        if (capturing) S3D_CK(cudaEventRecordWithFlags(ev, st, cudaEventRecordExternal));
        else S3D_CK(cudaEventRecord(ev, st));
        return 0;
    };
    // the first blur reads the caller's image: always launched directly
    if (!from_seed && run_level(0, -1)) return -1;

    auto body = [&]() -> int {
    // Phase A, the critical chain: every octave up to the level that seeds the next one.
    for (int o = 0; o < noct; o++) {
        if (o > 0 && !serial) S3D_CK(cudaStreamWaitEvent(stream_of(o), ctx->ev_seed[o - 1], 0));
        for (int s = 0; s <= ds_level && s < nlev - 1; s++)
            if (run_level(o, s)) return -1;
        if (o == 0 && early_levels > 0 && run_extrema(0, 0, early_levels)) return -1;
    }
    // Phase B, the bulk: remaining levels and the extrema scan of each octave. Octave 0's
    // one-wave kernels would occupy every SM while octave 1 still has to produce the seed of
    // octave 2; instead of waiting for that seed (SIFT3D_CUDA_SEEDWAIT >= 0: wait for the seed
    // event of that octave) its first bulk level starts at once on 148 - 40 SMs and the later
    // ones on 148 - 14 (sweeps in DESIGN.md section 3.5).
    static const int wait_env = getenv("SIFT3D_CUDA_SEEDWAIT") ? atoi(getenv("SIFT3D_CUDA_SEEDWAIT")) : -1;
    {
        int wi = wait_env < noct - 2 ? wait_env : noct - 2;
        if (wi < 0) wi = 0;
        if (noct > 1 && !serial && wait_env >= 0) S3D_CK(cudaStreamWaitEvent(st, ctx->ev_seed[wi], 0));
    }
    // Octave 0's remaining levels share the device with the other octaves' streams: leave
    // them some SMs (see set_fused_sm_reserve).
    if (noct > 1 && !serial) set_fused_sm_reserve(reserve_env);
    // The bulk levels of octave o >= 1 are not on the seed chain; held back until the chain is
    // `bulk_wait` octaves further on they do not take the reserved SMs from it.
    static const int bulk_wait = getenv("SIFT3D_CUDA_BULKWAIT") ? atoi(getenv("SIFT3D_CUDA_BULKWAIT")) : 0;
    const int last_seed_ev = (tiny_from >= 0 ? tiny_from : noct_plan) - 1 < noct - 2
                                 ? (tiny_from >= 0 ? tiny_from : noct_plan) - 1 : noct - 2;
    for (int o = 0; o < noct; o++) {
        const int sfirst = 0;
        int rc = 0;
        if (o >= 1 && bulk_wait > 0 && !serial && !slab && !(tiny_from >= 0 && o >= tiny_from)) {
            const int we = o + bulk_wait < last_seed_ev ? o + bulk_wait : last_seed_ev;
            if (we > o) S3D_CK(cudaStreamWaitEvent(stream_of(o), ctx->ev_seed[we], 0));
        }
        for (int s = (ds_level + 1 > sfirst ? ds_level + 1 : sfirst); s < nlev - 1 && !rc; s++) {
            // the first bulk level of octave 0 may start while octave 1 still works on its seed
            // chain (wait_env < 0): it then leaves more SMs free than the later ones
            if (o == 0 && noct > 1 && !serial)
                set_fused_sm_reserve(s == ds_level + 1 ? reserve_first : reserve_env);
            rc = run_level(o, s);
        }
        if (o == 0) set_fused_sm_reserve(0);
        if (rc) return -1;
        if (o == 0 && mark(ctx->ev[EV_PYR])) return -1;
    }
    if (slab) {  // DoG maxima are over whole levels
        if (host_sync()) return -1;
        if (cb->reduce_max(cb->user, (float *) ctx->scalars.p,
                           (int) (ctx->scalars.bytes / sizeof(float))))
            return -1;
    }
    for (int o = 0; o < noct; o++) {
        cudaStream_t so = stream_of(o);
        if (run_extrema(o, o == 0 ? early_levels : 0, kp)) return -1;
        // Scan order (o, s, z, y, x) and assign_orientations (sift.c:1109-1143) per octave, on
        // the octave's own stream as soon as its extrema exist: candidate counts are read on
        // the device and results go to the octave's own region, so no octave waits for another.
        // (At 256^3 three quarters of the candidates are in octaves >= 1, which finish long
        // before octave 0 -- orienting everything after the last octave cost 0.2 ms.)
        if (o == 0 && mark(ctx->ev[EV_EXTREMA])) return -1;
        // The octaves of the tiny chain share a stream and are the tail of detect: their
        // candidates are ordered and oriented by ONE launch each, after the last of them has
        // been scanned (two launches of ~10 / ~27 us per octave otherwise).
        const bool tiny_group = tiny_from >= 0 && o >= tiny_from && !slab;
        if (tiny_group && o + 1 < noct) continue;
        const int ob = tiny_group ? tiny_from : o;
        if (launch_order_all(so, (const Cand *) ctx->cand_raw.p, (Cand *) ctx->cand_sorted.p,
                             ctx->cand_base.data(), ctx->cand_cap.data(), noct,
                             (const int *) ctx->counts.p, &ctx->launches, ob, o + 1, &ctx->order_scratch))
            return -1;
        if (o == 0 && mark(ctx->ev[EV_ORDER])) return -1;
        if (launch_orient_all(so, (const Cand *) ctx->cand_sorted.p, (const OrientPlan *) ctx->orient_plan.p,
                              (const int *) ctx->counts.p, (int *) ctx->work.p + ob, corner_thresh,
                              ctx->kps_dev(), &ctx->launches, ob, o + 1,
                              ctx->kps_host ? ctx->kflags_host : nullptr, shared_device))
            return -1;
        if (so != st)
            for (int od = ob; od <= o; od++) S3D_CK(cudaEventRecord(ctx->ev_done[od], so));
    }
    if (!serial)
        for (int o = 1; o < noct; o++) S3D_CK(cudaStreamWaitEvent(st, ctx->ev_done[o], 0));
    S3D_CK(cudaMemcpyAsync(ctx->h_counts, ctx->counts.p, sizeof(int) * noct,
                           cudaMemcpyDeviceToHost, st));
    if (mark(ctx->ev[EV_ORIENT])) return -1;
    if (ctx->kps_host && mark(ctx->ev[EV_D2H])) return -1;  // results are already in host memory
    return 0;
    };  // body

    if (replay) {
        S3D_CK(cudaGraphLaunch(ctx->graph_exec, st));
        ctx->launches += ctx->graph_launches;
    } else if (capturing) {
        const uint64_t l0 = ctx->launches;
        S3D_CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed));
        const int rc = body();
        cudaGraph_t g = nullptr;
        const cudaError_t e = cudaStreamEndCapture(st, &g);
        set_fused_sm_reserve(0);
        if (rc != 0 || e != cudaSuccess || g == nullptr) {
            S3D_ERR("sift3d_cuda_detect: graph capture failed (%s); running without a graph\n",
                    cudaGetErrorString(e));
            cudaGetLastError();
            if (g) cudaGraphDestroy(g);
            ctx->use_graph = false;
            ctx->launches = l0;
            capturing = false;
            if (body()) return -1;
        } else {
            cudaGraphExec_t ge = nullptr;
            // (the octave streams' priorities are part of the schedule: keep them per node)
            const cudaError_t ei = cudaGraphInstantiateWithFlags(&ge, g, cudaGraphInstantiateFlagUseNodePriority);
            cudaGraphDestroy(g);
            if (ei != cudaSuccess || ge == nullptr) {
                S3D_ERR("sift3d_cuda_detect: cudaGraphInstantiate failed (%s); running without a graph\n",
                        cudaGetErrorString(ei));
                cudaGetLastError();
                ctx->use_graph = false;
                ctx->launches = l0;
                capturing = false;
                if (body()) return -1;
            } else {
                ctx->graph_exec = ge;
                ctx->graph_launches = ctx->launches - l0;
                S3D_CK(cudaGraphLaunch(ge, st));
            }
        }
    } else if (body()) {
        return -1;
    }
    if (host_wait(ctx, st)) return -1;
    int total = 0;
    bool overflow = false;
    for (int o = 0; o < noct; o++) {
        if (ctx->h_counts[o] > ctx->cand_cap[o]) overflow = true;
        total += ctx->h_counts[o];
    }
    if (overflow) {
        // An unusually busy image (e.g. noise-free synthetic blobs): the extrema kernel kept
        // counting past the end of a region without writing. Size the regions for what was
        // counted and run the detection again (a slab cannot: its peers would not follow).
        for (int o = 0; o < noct; o++)
            if (ctx->h_counts[o] > ctx->cand_cap[o]) {
                if ((slab != nullptr && !cb->in_stream) || attempt >= 2) {
                    S3D_ERR("sift3d_cuda_detect: octave %d produced %d extrema, capacity %d\n", o,
                            ctx->h_counts[o], ctx->cand_cap[o]);
                    return -1;
                }
                ctx->cand_min[o] = ctx->h_counts[o] + ctx->h_counts[o] / 4 + 1024;
            }
        if (plan_candidates(ctx)) return -1;
        // a slab cannot run again on its own (its peers would not follow): the regions are
        // resized and the caller (sift3d_cuda_multi_detect) repeats the detect on every slab
        if (slab != nullptr) return SIFT3D_CUDA_SLAB_RETRY;
        return detect_impl(ctx, image, image_on_device, units, peak_thresh, corner_thresh,
                           num_candidates, from_seed, slab, cb, attempt + 1);
    }
    ctx->total_candidates = total;
    ctx->have_pyramid = true;
    if (ctx->kps_host) detect_timings(ctx);
    if (num_candidates) *num_candidates = total;
    return 0;
}

int sift3d_cuda_detect(sift3d_cuda_ctx *ctx, const float *image, int image_on_device,
                       const double units[3], double peak_thresh, double corner_thresh,
                       int *num_candidates) {
    return detect_impl(ctx, image, image_on_device, units, peak_thresh, corner_thresh,
                       num_candidates, false, nullptr, nullptr);
}

int sift3d_cuda_detect_slab(sift3d_cuda_ctx *ctx, const float *image, int image_on_device,
                            const double units[3], double peak_thresh, double corner_thresh,
                            const sift3d_cuda_slab_desc *slab,
                            const sift3d_cuda_slab_callbacks *cb, int *num_candidates) {
    if (!slab || !cb) {
        S3D_ERR("sift3d_cuda_detect_slab: NULL slab description\n");
        return -1;
    }
    return detect_impl(ctx, image, image_on_device, units, peak_thresh, corner_thresh,
                       num_candidates, false, slab, cb);
}

int sift3d_cuda_detect_from_seed(sift3d_cuda_ctx *ctx, const float *seed, int seed_on_device,
                                 const double units[3], double peak_thresh,
                                 double corner_thresh, int *num_candidates) {
    return detect_impl(ctx, seed, seed_on_device, units, peak_thresh, corner_thresh,
                       num_candidates, true, nullptr, nullptr);
}

int sift3d_cuda_set_exact(sift3d_cuda_ctx *ctx, int exact) {
    if (!ctx) return -1;
    ctx->exact = exact != 0;
    return 0;
}

int sift3d_cuda_get_exact(sift3d_cuda_ctx *ctx) { return ctx ? (ctx->exact ? 1 : 0) : -1; }

int sift3d_cuda_set_base_octave(sift3d_cuda_ctx *ctx, int base_octave) {
    if (!ctx || base_octave < 0) return -1;
    ctx->base_octave = base_octave;
    return 0;
}

float *sift3d_cuda_level_ptr(sift3d_cuda_ctx *ctx, int o, int s) {
    if (!ctx || !ctx->planned || o < 0 || o >= ctx->desc.num_octaves || s < -1 ||
        s > ctx->nlev - 2)
        return nullptr;
    return ctx->level(o, s).d;
}

static void detect_timings(sift3d_cuda_ctx *ctx) {
    float ms;
    sift3d_cuda_timings &t = ctx->tm;
    cudaEventElapsedTime(&ms, ctx->ev[EV_START], ctx->ev[EV_H2D]); t.h2d = ms;
    cudaEventElapsedTime(&ms, ctx->ev[EV_H2D], ctx->ev[EV_SCALE]); t.scale = ms;
    cudaEventElapsedTime(&ms, ctx->ev[EV_SCALE], ctx->ev[EV_PYR]); t.pyramid = ms;
    t.dogmax = 0;
    cudaEventElapsedTime(&ms, ctx->ev[EV_PYR], ctx->ev[EV_EXTREMA]); t.extrema = ms;
    cudaEventElapsedTime(&ms, ctx->ev[EV_EXTREMA], ctx->ev[EV_ORDER]); t.order = ms;
    cudaEventElapsedTime(&ms, ctx->ev[EV_ORDER], ctx->ev[EV_ORIENT]); t.orient = ms;
    cudaEventElapsedTime(&ms, ctx->ev[EV_ORIENT], ctx->ev[EV_D2H]); t.d2h = ms;
    cudaEventElapsedTime(&ms, ctx->ev[EV_START], ctx->ev[EV_D2H]); t.total_detect = ms;
}

int sift3d_cuda_fetch_candidates(sift3d_cuda_ctx *ctx, sift3d_cuda_keypoint *out, int max) {
    if (!ctx || !ctx->have_pyramid || (!out && max > 0)) return -1;
    if (set_device(ctx)) return -1;
    int n = ctx->total_candidates < max ? ctx->total_candidates : max;
    {
        // results live in per-octave regions on the device; the caller gets them packed in
        // scan order (octave-major)
        int done = 0;
        for (size_t o = 0; o < ctx->cand_base.size() && done < n; o++) {
            int c = ctx->h_counts[o];
            if (c > n - done) c = n - done;
            if (c > 0 && ctx->kps_host)  // already in host memory (detect synchronised)
                memcpy(out + done, ctx->kps_host + ctx->cand_base[o], sizeof(sift3d_cuda_keypoint) * (size_t) c);
            else if (c > 0)
                S3D_CK(cudaMemcpyAsync(out + done,
                                       (const sift3d_cuda_keypoint *) ctx->kps.p + ctx->cand_base[o],
                                       sizeof(sift3d_cuda_keypoint) * (size_t) c, cudaMemcpyDeviceToHost,
                                       ctx->stream));
            done += c;
        }
    }
    if (ctx->kps_host) return n;  // timings were closed by detect itself
    S3D_CK(cudaEventRecord(ctx->ev[EV_D2H], ctx->stream));
    S3D_CK(cudaStreamSynchronize(ctx->stream));
    detect_timings(ctx);
    return n;
}

int sift3d_cuda_candidates_view(sift3d_cuda_ctx *ctx, int octave_index,
                                const sift3d_cuda_keypoint **records) {
    if (!ctx || !ctx->have_pyramid || !records || !ctx->kps_host || octave_index < 0 ||
        (size_t) octave_index >= ctx->cand_base.size())
        return -1;
    *records = ctx->kps_host + ctx->cand_base[octave_index];
    return ctx->h_counts[octave_index];
}

int sift3d_cuda_candidates_flags(sift3d_cuda_ctx *ctx, int octave_index,
                                 const sift3d_cuda_candidate_flag **flags) {
    if (!ctx || !ctx->have_pyramid || !flags || !ctx->kps_host || !ctx->kflags_host || octave_index < 0 ||
        (size_t) octave_index >= ctx->cand_base.size())
        return -1;
    *flags = ctx->kflags_host + ctx->cand_base[octave_index];
    return ctx->h_counts[octave_index];
}

static int describe_impl(sift3d_cuda_ctx *ctx, const sift3d_cuda_keypoint *keys, int num,
                         float *desc, int desc_on_device, const int *rows);

int sift3d_cuda_describe(sift3d_cuda_ctx *ctx, const sift3d_cuda_keypoint *keys, int num,
                         float *desc, int desc_on_device) {
    return describe_impl(ctx, keys, num, desc, desc_on_device, nullptr);
}

int sift3d_cuda_describe_rows(sift3d_cuda_ctx *ctx, const sift3d_cuda_keypoint *keys, int num,
                              float *desc, int desc_on_device, const int *rows) {
    if (!rows) {
        S3D_ERR("sift3d_cuda_describe_rows: NULL row table\n");
        return -1;
    }
    return describe_impl(ctx, keys, num, desc, desc_on_device, rows);
}

static int describe_impl(sift3d_cuda_ctx *ctx, const sift3d_cuda_keypoint *keys, int num,
                         float *desc, int desc_on_device, const int *rows) {
    if (!ctx || !ctx->have_pyramid) {
        S3D_ERR("sift3d_cuda_describe: no Gaussian pyramid is available; run detect first\n");
        return -1;
    }
    if (num < 1 || !keys || !desc) {
        S3D_ERR("sift3d_cuda_describe: invalid number of keypoints: %d\n", num);
        return -1;
    }
    if (set_device(ctx)) return -1;
    cudaStream_t st = ctx->stream;
    const int noct = ctx->desc.num_octaves, kp = ctx->desc.num_kp_levels;
    // The job list is staged in page-locked memory and pulled by a kernel (zero-copy reads)
    // instead of a copy-engine transfer: a host-to-device copy would queue behind an image
    // upload that is in flight for the next volume (sift3d_cuda_upload_begin) and stall the
    // descriptor kernel for the length of that upload.
    if (ctx->h_jobs_cap < (size_t) num) {
        if (ctx->h_jobs) cudaFreeHost(ctx->h_jobs);
        ctx->h_jobs = nullptr;
        ctx->h_jobs_cap = 0;
        const size_t cap = (size_t) num * 2 + 256;
        S3D_CK(cudaMallocHost((void **) &ctx->h_jobs, cap * sizeof(DescJob)));
        ctx->h_jobs_cap = cap;
    }
    // largest windows first: the grid drains in launch order, so the long CTAs start early.
    // The order is a stable sort on the row bound; only (index, rows) pairs are moved and the
    // 96-byte jobs are written once, in their final order, into the page-locked list.
    std::vector<std::pair<int, int>> ord((size_t) num);  // (rows, index)
    for (int i = 0; i < num; i++) {
        const sift3d_cuda_keypoint &k = keys[i];
        // any Gaussian level may be addressed (the reference indexes gpyr by key->o, key->s)
        const int ko = k.o - ctx->base_octave;
        if (ko < 0 || ko >= noct || k.s < -1 || k.s > kp + 1) {
            S3D_ERR("sift3d_cuda_describe: keypoint %d has invalid octave/level (%d, %d)\n", i,
                    k.o, k.s);
            return -1;
        }
        const Level &L = ctx->level(ko, k.s);
        DescJob J;
        J.ny = L.ny; J.nz = L.nz;
        J.uy = (float) L.uy; J.uz = (float) L.uz;
        J.sigma = (float) (k.sd * 7.071067812);
        ord[(size_t) i] = std::make_pair(desc_job_rows(J), i);
    }
    {
        // The row bound takes one value per (octave, level): a stable counting sort over the few
        // distinct values (descending) instead of n log n comparisons (30 us for 1444 jobs).
        int vals[64], cnt[64], nv = 0;
        bool few = true;
        for (int i = 0; i < num && few; i++) {
            int k = 0;
            while (k < nv && vals[k] != ord[(size_t) i].first) k++;
            if (k == nv) {
                if (nv == 64) { few = false; break; }
                vals[nv] = ord[(size_t) i].first;
                cnt[nv++] = 0;
            }
            cnt[k]++;
        }
        if (few) {
            int idx[64], start[64];
            for (int k = 0; k < nv; k++) idx[k] = k;
            std::sort(idx, idx + nv, [&](int a, int b) { return vals[a] > vals[b]; });
            int acc = 0;
            for (int r = 0; r < nv; r++) { start[idx[r]] = acc; acc += cnt[idx[r]]; }
            std::vector<std::pair<int, int>> tmp((size_t) num);
            int last_k = 0;
            for (int i = 0; i < num; i++) {
                const int v = ord[(size_t) i].first;
                int k = last_k;
                if (vals[k] != v) { k = 0; while (vals[k] != v) k++; last_k = k; }
                tmp[(size_t) start[k]++] = ord[(size_t) i];
            }
            ord.swap(tmp);
        } else {
            std::stable_sort(ord.begin(), ord.end(), [](const std::pair<int, int> &a, const std::pair<int, int> &b) {
                return a.first > b.first;
            });
        }
    }
    DescJob *jobs = ctx->h_jobs;
    for (int r = 0; r < num; r++) {
        const int i = ord[(size_t) r].second;
        const sift3d_cuda_keypoint &k = keys[i];
        const Level &L = ctx->level(k.o - ctx->base_octave, k.s);
        DescJob &J = jobs[r];
        J.img = L.d;
        J.nx = L.nx; J.ny = L.ny; J.nz = L.nz;
        J.ux = (float) L.ux; J.uy = (float) L.uy; J.uz = (float) L.uz;
        J.cx = (float) k.x; J.cy = (float) k.y; J.cz = (float) k.z;  // sift.c:1474-1476
        J.sigma = (float) (k.sd * 7.071067812);                        // sift.c:43,1453
        for (int j = 0; j < 9; j++) J.R[j] = k.R[j];
        J.out_index = rows ? rows[i] : i;
        J.max_rows = ord[(size_t) r].first;
    }
    const int max_rows = jobs[0].max_rows;
    if (ctx->desc_jobs.ensure(sizeof(DescJob) * (size_t) num)) return -1;
    float *d_out = desc;
    // Page-locked result buffer (the descriptor store of libsift3D.so is one): every CTA
    // writes its 768 floats straight into host memory as it finishes (posted PCIe writes under
    // the kernel) instead of a 4.4 MB device-to-host copy after it (86 us at 256^3).
    bool direct = false;
    if (!desc_on_device) {
        static const bool no_direct = getenv("SIFT3D_CUDA_NO_DIRECT_DESC") != nullptr;
        cudaPointerAttributes at;
        if (!no_direct && cudaPointerGetAttributes(&at, desc) == cudaSuccess &&
            at.type == cudaMemoryTypeHost && at.devicePointer != nullptr) {
            d_out = (float *) at.devicePointer;
            direct = true;
        } else {
            cudaGetLastError();
            if (rows) {
                S3D_ERR("sift3d_cuda_describe_rows: the descriptor buffer must be device-accessible "
                        "(device memory or page-locked host memory)\n");
                return -1;
            }
            if (ctx->desc_out.ensure(sizeof(float) * SIFT3D_CUDA_DESC_NUMEL * (size_t) num)) return -1;
            d_out = (float *) ctx->desc_out.p;
        }
    }
    S3D_CK(cudaEventRecord(ctx->dv[DV_START], st));
    if (launch_pull_words(st, ctx->h_jobs, ctx->desc_jobs.p, sizeof(DescJob) * (size_t) num / 4,
                          &ctx->launches))
        return -1;
    S3D_CK(cudaEventRecord(ctx->dv[DV_H2D], st));
    if (launch_descriptors(st, (const DescJob *) ctx->desc_jobs.p, num, max_rows, d_out, &ctx->launches))
        return -1;
    S3D_CK(cudaEventRecord(ctx->dv[DV_KERNEL], st));
    if (!desc_on_device && !direct)
        S3D_CK(cudaMemcpyAsync(desc, d_out, sizeof(float) * SIFT3D_CUDA_DESC_NUMEL * (size_t) num,
                               cudaMemcpyDeviceToHost, st));
    S3D_CK(cudaEventRecord(ctx->dv[DV_D2H], st));
    if (host_wait(ctx, st)) return -1;
    float ms;
    cudaEventElapsedTime(&ms, ctx->dv[DV_START], ctx->dv[DV_H2D]); ctx->tm.desc_h2d = ms;
    cudaEventElapsedTime(&ms, ctx->dv[DV_H2D], ctx->dv[DV_KERNEL]); ctx->tm.descriptors = ms;
    cudaEventElapsedTime(&ms, ctx->dv[DV_KERNEL], ctx->dv[DV_D2H]); ctx->tm.desc_d2h = ms;
    cudaEventElapsedTime(&ms, ctx->dv[DV_START], ctx->dv[DV_D2H]); ctx->tm.total_describe = ms;
    return 0;
}

int sift3d_cuda_level_info(sift3d_cuda_ctx *ctx, int o, int s, int dims[3], double units[3],
                           double *scale) {
    if (!ctx || !ctx->planned || o < 0 || o >= ctx->desc.num_octaves || s < -1 || s > ctx->nlev - 2)
        return -1;
    const Level &L = ctx->level(o, s);
    if (dims) { dims[0] = L.nx; dims[1] = L.ny; dims[2] = L.nz; }
    if (units) { units[0] = L.ux; units[1] = L.uy; units[2] = L.uz; }
    if (scale) *scale = L.scale;
    return 0;
}

int sift3d_cuda_fetch_level(sift3d_cuda_ctx *ctx, int o, int s, float *host_out) {
    if (!ctx || !ctx->have_pyramid || !host_out || o < 0 || o >= ctx->desc.num_octaves || s < -1 ||
        s > ctx->nlev - 2)
        return -1;
    if (set_device(ctx)) return -1;
    const Level &L = ctx->level(o, s);
    S3D_CK(cudaMemcpyAsync(host_out, L.d, L.nvox() * sizeof(float), cudaMemcpyDeviceToHost,
                           ctx->stream));
    S3D_CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int sift3d_cuda_dogmax(sift3d_cuda_ctx *ctx, int o, int s, float *out) {
    if (!ctx || !ctx->have_pyramid || !out || o < 0 || o >= ctx->desc.num_octaves || s < 0 ||
        s >= ctx->desc.num_kp_levels)
        return -1;
    if (set_device(ctx)) return -1;
    S3D_CK(cudaMemcpyAsync(out, ctx->d_scalar(1 + o * SIFT3D_CUDA_MAX_FILTERS + s), sizeof(float),
                           cudaMemcpyDeviceToHost, ctx->stream));
    S3D_CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int sift3d_cuda_blur(sift3d_cuda_ctx *ctx, const float *host_src, float *host_dst, int nx, int ny,
                     int nz, const double units[3], const sift3d_cuda_filter *f,
                     int force_generic) {
    if (!ctx || !host_src || !host_dst || !units || !f || nx < 2 || ny < 2 || nz < 2 ||
        f->width < 1 || f->width > SIFT3D_CUDA_MAX_TAPS || !(f->width & 1))
        return -1;
    if (set_device(ctx)) return -1;
    const size_t n = (size_t) nx * ny * nz;
    DevBuf a, b, c, d, t;
    int ret = -1;
    if (a.ensure(n * 4) || b.ensure(n * 4) || c.ensure(n * 4) || d.ensure(n * 4) ||
        t.ensure(sizeof(float) * SIFT3D_CUDA_MAX_TAPS))
        goto done;
    if (cudaMemcpyAsync(a.p, host_src, n * 4, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess ||
        cudaMemcpyAsync(t.p, f->taps, sizeof(float) * f->width, cudaMemcpyHostToDevice,
                        ctx->stream) != cudaSuccess)
        goto done;
    if (launch_blur(ctx->stream, (const float *) a.p, (float *) b.p, (float *) c.p, (float *) d.p,
                    nx, ny, nz, units, *f, (const float *) t.p, nullptr, nullptr, nullptr, nullptr,
                    force_generic != 0, &ctx->launches, 0, -1, 0, ctx->exact))
        goto done;
    if (cudaMemcpyAsync(host_dst, b.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
        cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
        S3D_ERR("sift3d_cuda_blur: %s\n", cudaGetErrorString(cudaGetLastError()));
        goto done;
    }
    ret = 0;
done:
    a.release(); b.release(); c.release(); d.release(); t.release();
    return ret;
}

int sift3d_cuda_get_timings(sift3d_cuda_ctx *ctx, sift3d_cuda_timings *t) {
    if (!ctx || !t) return -1;
    *t = ctx->tm;
    return 0;
}

uint64_t sift3d_cuda_launch_count(sift3d_cuda_ctx *ctx) { return ctx ? ctx->launches : 0; }

void *sift3d_cuda_device_upload(int device, const float *host, size_t count) {
    void *p = nullptr;
    if (!host || count == 0 || cudaSetDevice(device) != cudaSuccess ||
        cudaMalloc(&p, count * sizeof(float)) != cudaSuccess) {
        S3D_ERR("sift3d_cuda_device_upload: %s\n", cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    if (cudaMemcpy(p, host, count * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) {
        S3D_ERR("sift3d_cuda_device_upload: %s\n", cudaGetErrorString(cudaGetLastError()));
        cudaFree(p);
        return nullptr;
    }
    return p;
}

void sift3d_cuda_device_free(int device, void *dptr) {
    if (!dptr) return;
    cudaSetDevice(device);
    cudaFree(dptr);
}

struct sift3d_cuda_upload {
    int device = 0;
    void *d = nullptr;
    size_t bytes = 0;
    cudaStream_t stream = nullptr;  // the device's upload stream (shared, not owned)
    cudaEvent_t done = nullptr;
};

// One upload stream per device, shared by every image: host-to-device copies run one after
// another on the copy engine anyway, and a stream per image (16 of them under bench.py's four
// pipelines, next to 7 streams per detector) made the streams share the device's hardware
// queues -- an upload then sat behind another pipeline's kernels (seen as a bimodal e2e time).
static cudaStream_t upload_stream(int device) {
    static std::mutex mu;
    static cudaStream_t streams[SIFT3D_CUDA_MULTI_MAX_DEVICES] = {};
    if (device < 0 || device >= SIFT3D_CUDA_MULTI_MAX_DEVICES) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    if (!streams[device]) {
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        if (cudaStreamCreateWithPriority(&streams[device], cudaStreamNonBlocking, hi) != cudaSuccess) {
            cudaGetLastError();
            streams[device] = nullptr;
        }
    }
    return streams[device];
}

// The library runs many streams per device (octave streams of every detector, uploads): ask
// for the maximum number of hardware queues unless the caller chose a value. Read by the
// driver when the CUDA context is created, so it is set when the library is loaded.
__attribute__((constructor)) static void s3d_more_hardware_queues() {
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", /* overwrite */ 0);
}

void sift3d_cuda_upload_free(sift3d_cuda_upload *up) {
    if (!up) return;
    cudaSetDevice(up->device);
    if (up->done) {
        cudaEventSynchronize(up->done);
        cudaEventDestroy(up->done);
    }
    if (up->d) cudaFreeAsync(up->d, up->stream);
    delete up;
}

sift3d_cuda_upload *sift3d_cuda_upload_begin(int device, const float *host, size_t count,
                                             sift3d_cuda_upload *reuse) {
    if (!host || count == 0) {
        S3D_ERR("sift3d_cuda_upload_begin: empty buffer\n");
        return nullptr;
    }
    sift3d_cuda_upload *up = reuse;
    if (up && up->device != device) {
        sift3d_cuda_upload_free(up);
        up = nullptr;
    }
    bool fresh = false;
    if (!up) {
        up = new (std::nothrow) sift3d_cuda_upload();
        if (!up) return nullptr;
        up->device = device;
        fresh = true;
    }
    const size_t bytes = count * sizeof(float);
    bool ok = cudaSetDevice(device) == cudaSuccess;
    if (ok && fresh) {
        up->stream = upload_stream(device);
        ok = up->stream != nullptr && cudaEventCreateWithFlags(&up->done, cudaEventDisableTiming) == cudaSuccess;
    }
    if (ok && up->bytes < bytes) {
        // stream-ordered allocation: cudaMalloc / cudaFree wait for an idle device, which other
        // pipelines of the process may not offer for a long time (seen as 60 ms stalls)
        if (up->d) {
            cudaFreeAsync(up->d, up->stream);
            up->d = nullptr;
            up->bytes = 0;
        }
        ok = cudaMallocAsync(&up->d, bytes, up->stream) == cudaSuccess;
        if (ok) up->bytes = bytes;
    }
    ok = ok && cudaMemcpyAsync(up->d, host, bytes, cudaMemcpyHostToDevice, up->stream) == cudaSuccess &&
         cudaEventRecord(up->done, up->stream) == cudaSuccess;
    if (!ok) {
        S3D_ERR("sift3d_cuda_upload_begin: %s\n", cudaGetErrorString(cudaGetLastError()));
        sift3d_cuda_upload_free(up);
        return nullptr;
    }
    return up;
}

const float *sift3d_cuda_upload_ptr(const sift3d_cuda_upload *up) {
    return up ? (const float *) up->d : nullptr;
}

int sift3d_cuda_upload_device(const sift3d_cuda_upload *up) { return up ? up->device : -1; }

int sift3d_cuda_upload_sync(sift3d_cuda_upload *up) {
    if (!up) return -1;
    S3D_CK(cudaSetDevice(up->device));
    S3D_CK(cudaEventSynchronize(up->done));
    return 0;
}

int sift3d_cuda_upload_wait(sift3d_cuda_ctx *ctx, sift3d_cuda_upload *up) {
    if (!ctx || !up || up->device != ctx->device) return -1;
    if (set_device(ctx)) return -1;
    S3D_CK(cudaStreamWaitEvent(ctx->stream, up->done, 0));
    return 0;
}

int sift3d_cuda_ctx_device(sift3d_cuda_ctx *ctx) { return ctx ? ctx->device : -1; }

void *sift3d_cuda_ctx_stream(sift3d_cuda_ctx *ctx) { return ctx ? (void *) ctx->stream : nullptr; }

int sift3d_cuda_mark(sift3d_cuda_ctx *ctx, int slot) {
    if (!ctx || slot < 0 || slot >= 8) return -1;
    if (set_device(ctx)) return -1;
    S3D_CK(cudaEventRecord(ctx->marks[slot], ctx->stream));
    return 0;
}

int sift3d_cuda_elapsed_ms(sift3d_cuda_ctx *ctx, int a, int b, float *ms) {
    if (!ctx || !ms || a < 0 || a >= 8 || b < 0 || b >= 8) return -1;
    if (set_device(ctx)) return -1;
    S3D_CK(cudaEventSynchronize(ctx->marks[b]));
    S3D_CK(cudaEventElapsedTime(ms, ctx->marks[a], ctx->marks[b]));
    return 0;
}

int sift3d_cuda_set_profiling(sift3d_cuda_ctx *ctx, int enabled) {
    if (!ctx) return -1;
    ctx->profiling = enabled != 0;
    return 0;
}

int sift3d_cuda_blur_ms(sift3d_cuda_ctx *ctx, int o, int s, float *ms, int *launches) {
    if (!ctx || !ctx->have_pyramid || !ctx->profiling || o < 0 || o >= ctx->desc.num_octaves ||
        s < -1 || s > ctx->nlev - 2 || (s == -1 && o != 0))
        return -1;
    if (set_device(ctx)) return -1;
    const size_t li = (size_t) o * ctx->nlev + (s + 1);
    S3D_CK(cudaEventSynchronize(ctx->blur_ev[2 * li + 1]));
    if (ms) S3D_CK(cudaEventElapsedTime(ms, ctx->blur_ev[2 * li], ctx->blur_ev[2 * li + 1]));
    if (launches) *launches = ctx->blur_launches[li];
    return 0;
}

int sift3d_cuda_level_span_ms(sift3d_cuda_ctx *ctx, int o, int s, float *begin_ms, float *end_ms) {
    if (!ctx || !ctx->have_pyramid || !ctx->profiling || o < 0 || o >= ctx->desc.num_octaves ||
        s < -1 || s > ctx->nlev - 2 || (s == -1 && o != 0))
        return -1;
    if (set_device(ctx)) return -1;
    const size_t li = (size_t) o * ctx->nlev + (s + 1);
    S3D_CK(cudaEventSynchronize(ctx->blur_ev[2 * li + 1]));
    if (begin_ms) S3D_CK(cudaEventElapsedTime(begin_ms, ctx->ev[EV_START], ctx->blur_ev[2 * li]));
    if (end_ms) S3D_CK(cudaEventElapsedTime(end_ms, ctx->ev[EV_START], ctx->blur_ev[2 * li + 1]));
    return 0;
}

int sift3d_cuda_synchronize(sift3d_cuda_ctx *ctx) {
    if (!ctx) return -1;
    if (set_device(ctx)) return -1;
    S3D_CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

}  // extern "C"
