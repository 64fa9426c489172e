#!/usr/bin/env python3
"""bench.py -- voxels/s of SIFT3D detect + describe on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the hot path (sift3d_detect_keypoints + sift3d_extract_descriptors)
over a batch of P synthetic 256^3 "MRI-like" volumes (BASELINE.json configs[1]) per GPU: P
independent pipelines ($SIFT3D_BENCH_PIPELINES, default 4: detector + stream + pyramid each)
driven by the C library's batch call, one volume per pipeline and step. With N > 1 every rank
owns one GPU and its own volumes (configs[3]: whole volumes per GPU, no data-path collective),
so scaling is "weak" and `value` = N * P * 256^3 * K / max-over-ranks time.

  value        inputs resident in HBM before the timed region (sift3d_image_make_resident),
               timed with CUDA events on the library's stream, device idle at both marks
  e2e          the same batch call with HOST (page-locked) image buffers: every 64 MiB upload,
               candidate download and descriptor download is inside the timed region
  single_pipeline / single_volume_latency_ms   one pipeline alone (resident) / one volume at a
               time from the host: what a caller with one detector and no batch sees
  roofline     the blur (octave-0 Gaussian levels): 8 B per voxel per level / CUDA-event time
  cpu_baseline the unmodified reference (oracle/_ref) on this box's host cores, bounded sample

The reference arm (--impl reference) times the reference's own CPU implementation.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

# many streams per device (octave streams of every pipeline, uploads): all the hardware queues the
# driver offers; read at context creation, so set before anything initialises CUDA
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

DIM = 256
NVOX = DIM ** 3
NUM_INPUTS = 4  # rotating inputs: 4 x 64 MiB > 126 MB L2
BLOBS = NVOX // 4096
FLAGS = 3  # bias field + noise ("MRI-like")
SEED = 88172645463325252
WORKLOAD = "C2_mri_256: 256^3 float32 blobs x bias + N(0,0.02) noise, units (1,1,1), default params"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def make_volume(i, out=None):
    from sift3d_b200 import synth_volume
    return synth_volume(DIM, DIM, DIM, num_blobs=BLOBS, seed=SEED + i, flags=FLAGS, out=out)


# ---------------------------------------------------------------------------------------
# CPU side: the reference's own implementation (oracle/_ref), or the restatement if the
# reference could not be compiled. Only this part of bench.py touches oracle/.
# ---------------------------------------------------------------------------------------
def cpu_oracle():
    from oracle import port, refstage
    if refstage.available():
        return "reference", refstage.RefRun
    if port.available():
        return "port", port.PortRun
    raise RuntimeError("no CPU oracle built: run `make -C oracle`")


class CpuRunner:
    """The reference's detect + describe on the FULL 256^3 workload volume, one detector reused
    across steps (so the reference re-plans its pyramids only once, sift.c:651-656, exactly
    like the product arm's detector)."""

    def __init__(self, Run):
        self.vol = make_volume(0)
        self.run = Run(self.vol, (1.0, 1.0, 1.0))

    def step(self):
        """detect + describe once on the CPU; returns (seconds, keypoints)."""
        run = self.run
        t0 = time.perf_counter()
        keys = run.detect_public() if hasattr(run, "detect_public") else run.detect()
        if len(keys):
            run.extract()
        return time.perf_counter() - t0, len(keys)

    def close(self):
        self.run.close()


def cpu_threads():
    n = os.environ.get("OMP_NUM_THREADS")
    return int(n) if n else (os.cpu_count() or 1)


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 into every rank when the variable is unset; the CPU
    legs run on rank 0 alone and are meant to use every host thread (the reference's OpenMP
    path). Must run before the oracle library (libgomp) is loaded."""
    if os.environ.get("TORCHELASTIC_RUN_ID") is not None and os.environ.get("OMP_NUM_THREADS") == "1":
        try:
            n = len(os.sched_getaffinity(0))
        except AttributeError:
            n = os.cpu_count() or 1
        os.environ["OMP_NUM_THREADS"] = str(n)


def run_reference_arm(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    use_all_host_threads()
    os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
    kind, Run = cpu_oracle()
    cpu = CpuRunner(Run)
    # one full 256^3 step costs several seconds on the host: one warm-up step (it also makes the
    # reference allocate its pyramids) keeps the whole arm within a few minutes
    warm = min(max(args.warmup, 0), 1)
    for _ in range(warm):
        cpu.step()
    t0 = time.perf_counter()
    nkeys = 0
    for _ in range(args.steps):
        _, nkeys = cpu.step()
    dt = time.perf_counter() - t0
    cpu.close()
    value = NVOX * args.steps / dt
    sample_desc = (f"the full 256^3 workload volume per step ({nkeys} keypoints), one detector "
                   f"reused across steps, all host threads")
    line = {
        "impl": "reference", "metric": "voxels/sec (256^3 detect+describe)", "value": value,
        "unit": "voxels/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": warm,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "volumes_per_step": 1, "sample": sample_desc},
        "cpu_baseline": {"value": value, "unit": "voxels/s", "cores": cpu_threads(), "kind": kind,
                         "sample": sample_desc},
        "e2e": {"value": value, "unit": "voxels/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_product_arm(args):
    import sift3d_b200 as s3

    rank, world, local = dist_env()
    if rank == 0:
        use_all_host_threads()  # the workload generator (OpenMP) and the CPU leg run on rank 0
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if s3.cuda_device_count() <= local:
        raise SystemExit("bench.py: no CUDA device for this rank; the product has no CPU path")

    # P independent pipelines (detector + stream + pyramid) per GPU, each on a host thread of the
    # C library (sift3d_detect_describe_batch: volume v on detector v % P). One pipeline leaves
    # the device idle in its launch gaps, in the tail of the descriptor kernel and -- with host
    # inputs -- during its upload; the kernels of the other pipelines fill those holes.
    P = max(1, int(os.environ.get("SIFT3D_BENCH_PIPELINES", "4")))
    dets = [s3.Detector(local) for _ in range(P)]
    det = dets[0]
    mode = "exact (0 ULP vs the reference)" if det.exact() else \
        "fast (one FFMA per tap; voxels within 1e-5 of the level's max-abs)"
    # every pipeline rotates through its own copies of the NUM_INPUTS distinct volumes (an image
    # object carries its upload state, so pipelines do not share them)
    pools = []
    for p_ in range(P):
        pool = []
        for i in range(NUM_INPUTS):
            im = s3.Image(DIM, DIM, DIM)
            if p_ == 0:
                make_volume(rank * NUM_INPUTS + i, out=im.array)
            else:
                np.copyto(im.array, pools[0][i].array)
            pool.append(im)
        pools.append(pool)
    images = pools[0]
    kps = [s3.KeypointStore() for _ in range(P)]
    dss = [s3.DescriptorStore() for _ in range(P)]
    kp, ds = kps[0], dss[0]

    def run_batch(nsteps, first=0):
        """nsteps steps = nsteps volumes on each of the P pipelines, one C call."""
        n = nsteps * P
        ims = [pools[v % P][(first + v // P) % NUM_INPUTS] for v in range(n)]
        s3.detect_describe_batch(dets, ims, kps=[kps[v % P] for v in range(n)],
                                 dss=[dss[v % P] for v in range(n)])

    def step(i):  # one volume on pipeline 0 alone (latency, profiling pass)
        det.detect(images[i % NUM_INPUTS], kp)
        if len(kp):
            det.extract(kp, ds)

    def barrier():
        if dist is not None:
            dist.barrier()
            import torch
            torch.cuda.synchronize()
        for d in dets:
            s3.libcuda().sift3d_cuda_synchronize(d.ctx)

    def timed(nsteps, slot_a, slot_b, batch, collect_blur=False):
        """K steps bracketed by barrier + synchronize; CUDA-event time, max over ranks. The
        marks are recorded on pipeline 0's stream with the device idle on both sides (every
        pipeline's last call has returned its results to the host)."""
        blur = {}
        barrier()
        l0 = sum(d.launch_count() for d in dets)
        t0 = time.perf_counter()
        det.mark(slot_a)
        if batch:
            run_batch(nsteps)
        else:
            for i in range(nsteps):
                step(i)
                if collect_blur:
                    for s in range(-1, 5):
                        ms, n = det.blur_ms(0, s)
                        acc = blur.setdefault(s, [0.0, n])
                        acc[0] += ms
        for d in dets[1:]:
            s3.libcuda().sift3d_cuda_synchronize(d.ctx)
        det.mark(slot_b)
        barrier()
        wall = time.perf_counter() - t0
        ms = det.elapsed_ms(slot_a, slot_b)
        launches = sum(d.launch_count() for d in dets) - l0
        if dist is not None:
            import torch
            t = torch.tensor([ms, wall * 1e3], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1]) / 1e3
        return ms, wall, launches, blur

    # ---- device-resident arm ----
    for pool in pools:
        for im in pool:
            im.make_resident(local)
    run_batch(max(args.warmup, 3))
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms, wall, launches, _ = timed(args.steps, 0, 1, True)
    clocks = sampler.stop() if rank == 0 else None
    ncand = len(det.candidates())
    nkeys = len(kp)
    # one pipeline alone: stage times and the per-level blur times (the roofline kernel) with
    # the per-blur event pairs switched on -- same inputs, not part of the headline timing
    for i in range(2):
        step(i)
    s_ms, _, _, _ = timed(args.steps, 6, 7, False)
    stage = det.timings()
    det.set_profiling(True)
    step(0)
    _, _, _, blur = timed(args.steps, 4, 5, False, collect_blur=True)
    det.set_profiling(False)

    # ---- end-to-end arm: host buffers, copies inside the timed region ----
    for pool in pools:
        for im in pool:
            im.make_resident(-1)
    for i in range(2):
        step(i)
    # (a) one volume at a time on one pipeline: upload, detect, describe, strictly in sequence
    l_ms, _, _, _ = timed(args.steps, 2, 3, False)
    e_stage = det.timings()
    # (b) the public batch call on host images: every volume is copied host->device (the worker
    # of a pipeline starts the upload of its next volume between detect and extract of the
    # current one, sift3d_image_prefetch) and its results device->host inside the timed region
    # (warm-up over every input of every pipeline: the first upload of an image allocates its
    # device buffer)
    run_batch(max(args.warmup, NUM_INPUTS))
    e_ms, e_wall, _, _ = timed(args.steps, 2, 3, True)
    # diagnostic ($SIFT3D_BENCH_E2E_REPEATS=n): the same measurement n more times
    e_more = [timed(args.steps, 2, 3, True)[0] / args.steps / P
              for _ in range(int(os.environ.get("SIFT3D_BENCH_E2E_REPEATS", "0")))]

    # ---- config 5 at N > 1: ONE 1024^3 volume z-slab sharded over the N GPUs of this box by
    # rank 0's process (sift3d_detect_describe_sharded: one host thread per device, halo
    # exchange over NVLink peer copies); the other ranks wait at the barrier
    c5 = None
    if dist is not None:
        for im in images:
            im.make_resident(-1)
        # the other ranks wait on the HOST (gloo): an NCCL barrier would keep a spinning kernel
        # on their GPUs, which rank 0's slabs would then have to time-slice with
        import torch
        host_group = dist.new_group(backend="gloo")
        torch.cuda.synchronize()
        dist.barrier(group=host_group)
        if rank == 0 and not os.environ.get("SIFT3D_BENCH_NO_C5"):
            try:
                c5 = run_c5_sharded(world)
            except Exception as e:  # noqa: BLE001 - reported in the line, the headline stands
                c5 = {"error": repr(e)}
        dist.barrier(group=host_group)

    # ---- the other single-GPU configurations of BASELINE.json, timed once each on rank 0
    # (N = 1 only): C1 = the CLI's call sequence on 128^3 (detect, keep the 100 "strongest",
    # describe; cli/kpSift3D.c:116-135), C3 = 512 x 512 x 300 with anisotropic units
    extras = None
    if rank == 0 and world == 1 and not os.environ.get("SIFT3D_BENCH_NO_EXTRAS"):
        # these are single-detector figures: the pipelines' detectors are released first (a
        # detector that shares its device launches smaller persistent grids)
        for d in dets:
            d.close()
        extras = run_other_configs(local)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    total_ms = sum(v[0] for v in blur.values()) / args.steps
    per_level = {str(s): {"ms": blur[s][0] / args.steps, "launches": blur[s][1],
                          "GBps": 8.0 * NVOX / (blur[s][0] / args.steps * 1e-3) / 1e9}
                 for s in sorted(blur)}
    achieved = 8.0 * NVOX * len(blur) / (total_ms * 1e-3) / 1e9
    # DRAM bytes per blur launch: not measurable in-run (ncu counters); the committed figure
    # of the ncu --set full capture of THIS kernel version, with its source, or null
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "r2_blur_traffic.json")
    if os.path.exists(tp):
        tj = json.load(open(tp))
        traffic, traffic_src = tj.get("dram_bytes_per_blur"), tj.get("source")
        # the figure belongs to one version of the kernel: stale once its source has changed
        ks = os.path.join(ROOT, tj.get("kernel_source", ""))
        if not tj.get("kernel_source_sha256") or not os.path.isfile(ks) or \
                hashlib.sha256(open(ks, "rb").read()).hexdigest() != tj["kernel_source_sha256"]:
            traffic, traffic_src = None, "profiles/r2_blur_traffic.json is older than the kernel source (re-run " \
                                         "tools/make_blur_traffic.py on a fresh ncu capture)"
    value = world * P * NVOX * args.steps / (ms * 1e-3)
    e2e = world * P * NVOX * args.steps / (e_ms * 1e-3)
    line = {
        "metric": "voxels/sec (256^3 detect+describe)", "value": value, "unit": "voxels/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "blur_arithmetic": mode, "volumes_per_gpu_per_step": P,
                   "parallelism": f"{P} independent pipelines (detector, stream, pyramid) per GPU x{world} "
                                  "GPUs, one volume per pipeline and step, no data-path collective",
                   "l2": f"per pipeline {NUM_INPUTS} rotating inputs (256 MiB) and a 0.9 GB pyramid rewritten "
                         f"every step, {P} pipelines: working set >> 126 MB L2",
                   "candidates": ncand, "keypoints": nkeys},
        "wall_ms_per_step": 1e3 * wall / args.steps,
        "ms_per_volume": ms / args.steps / P,
        "single_pipeline": {"ms_per_volume": s_ms / args.steps,
                            "voxels_per_s": world * NVOX * args.steps / (s_ms * 1e-3),
                            "stage_ms": stage},
        "stage_ms": stage,
        "e2e": {"value": e2e, "unit": "voxels/s", "ms_per_step": e_ms / args.steps,
                "ms_per_volume": e_ms / args.steps / P,
                "mode": f"sift3d_detect_describe_batch on page-locked HOST images, {P} pipelines per GPU: "
                        "every volume is uploaded and its results downloaded inside the timed region "
                        f"(K steps, {P} K uploads per GPU)",
                "single_volume_latency_ms": l_ms / args.steps,
                **({"repeats_ms_per_volume": e_more} if e_more else {}),
                "h2d_bytes_per_step": P * (NVOX * 4 + nkeys * 96),
                "d2h_bytes_per_step": P * (ncand * 88 + nkeys * 768 * 4 + 64 * 4),
                "stage_ms": e_stage},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"kernel": "blur (all launches producing one octave-0 Gaussian level; "
                               "6 levels per step)",
                     "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src, "timed": "CUDA events per blur in a separate pass of "
                     f"{args.steps} steps of ONE pipeline after the headline timing (the detect's own "
                     "concurrent schedule of streams; no second pipeline on the device)",
                     "algorithmic_bytes_per_launch": 8 * NVOX, "per_level": per_level},
        # SIFT3D_BENCH_NO_CPU=1: tuning runs only (the line is then incomplete by contract)
        # (rank 0 at N=1 only: at N>1 the reference arm carries the CPU number)
        "cpu_baseline": None if (os.environ.get("SIFT3D_BENCH_NO_CPU") or world > 1) else measure_cpu_baseline(),
    }
    if c5 is not None:
        line["c5_slab"] = c5
    if extras is not None:
        line["other_configs"] = extras
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_other_configs(device, reps=5):
    """Best-of-`reps` (C1: 3 x reps) host wall clock of the public call sequence, host image (upload inside)."""
    import sift3d_b200 as s3
    from sift3d_b200.synth import WORKLOADS, workload

    out = {}
    for name, limit in (("C1_blobs_128", 100), ("C3_ct_512x512x300", None)):
        (nx, ny, nz), units, _, _ = WORKLOADS[name]
        im = s3.Image(nx, ny, nz)
        workload(name, out=im.array)
        im.set_units(*units)
        det = s3.Detector(device)
        kp, ds = s3.KeypointStore(), s3.DescriptorStore()
        best, stage = None, None
        # (the small volume's upload time settles over the first repetitions: 0.3 -> 0.16 ms)
        for i in range((3 * reps if limit is not None else reps) + 1):
            t0 = time.perf_counter()
            det.detect(im, kp)
            nall = len(kp)
            if limit is not None:
                kp.sort_by_strength(limit)
            if len(kp):
                det.extract(kp, ds)
            dt = time.perf_counter() - t0
            if i > 0 and (best is None or dt < best):
                best, stage = dt, det.timings()
        out[name] = {"dims": [nx, ny, nz], "units": list(units), "e2e_ms": 1e3 * best,
                     "voxels_per_s": nx * ny * nz / best, "keypoints_detected": nall,
                     "keypoints_described": len(kp),
                     "stage_ms": {k: round(v, 4) for k, v in stage.items() if v}}
        det.close()
    return out


def run_c5_sharded(ngpus, dim=1024, reps=3):
    """BASELINE.json configs[4]: one dim^3 volume over `ngpus` GPUs (z-slabs) against the same
    volume on one GPU, through the public C API with the HOST image (uploads inside the time)."""
    import sift3d_b200 as s3
    from sift3d_b200 import synth_volume

    nvox = dim ** 3
    im = s3.Image(dim, dim, dim)
    synth_volume(dim, dim, dim, num_blobs=nvox // 4096, seed=SEED, flags=0, out=im.array)
    octaves = 4
    det = s3.Detector(0)
    kp, ds = s3.KeypointStore(), s3.DescriptorStore()
    best1 = None
    for _ in range(2):
        t0 = time.perf_counter()
        det.detect(im, kp)
        det.extract(kp, ds)
        dt = time.perf_counter() - t0
        best1 = dt if best1 is None else min(best1, dt)
    n1 = len(kp)
    det.close()
    det = s3.Detector(0)
    best, tm = None, None
    for i in range(reps + 1):  # the first run builds the per-device contexts
        t0 = time.perf_counter()
        det.detect_describe_sharded(im, list(range(ngpus)), octaves, kp=kp, ds=ds)
        dt = time.perf_counter() - t0
        if i > 0 and (best is None or dt < best):
            best, tm = dt, det.sharded_timings()
    out = {"workload": f"C5: one {dim}^3 float32 blob volume, units (1,1,1), z-slabs over {ngpus} GPUs "
                       f"of one process, {octaves} sharded octaves, host image (upload inside the time)",
           "n_gpus": ngpus, "e2e_ms": 1e3 * best, "value": nvox / best, "unit": "voxels/s",
           "single_gpu_e2e_ms": 1e3 * best1, "speedup_vs_single_gpu": best1 / best,
           "keypoints": len(kp), "keypoints_single_gpu": n1, "timings_ms": tm}
    det.close()
    return out


def measure_cpu_baseline():
    os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
    kind, Run = cpu_oracle()
    cpu = CpuRunner(Run)
    dt, nkeys = cpu.step()  # includes the reference's one-time pyramid allocation
    dt2, nkeys = cpu.step()
    cpu.close()
    best = min(dt, dt2)
    return {"value": NVOX / best, "unit": "voxels/s", "cores": cpu_threads(), "kind": kind,
            "sample": f"the full 256^3 workload volume, detect+describe ({nkeys} keypoints), "
                      f"best of 2 steps on one reused detector, {best:.2f} s"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    if args.steps < 1:
        raise SystemExit("--steps must be >= 1")
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_product_arm(args)


if __name__ == "__main__":
    main()
