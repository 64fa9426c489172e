"""One tall volume over several GPUs: z-slab sharding of the detect + describe path.

The reference has no multi-device path; this is the
decomposition of its octave loops (build_gpyr sift.c:662-711, detect_extrema
sift.c:735-871, assign_orientations sift.c:1109-1167, extract_descrip sift.c:1442-1536)
over z, arranged so that every rank computes exactly the values a single device would:

  * rank r owns the global planes [g0, g1) (multiples of 2^K, K = sharded octaves) and holds
    the slab [a0, a1) = owned planes + `halo` planes on each cut face;
  * a blur writes only the owned planes of a level (C ABI sift3d_cuda_detect_slab); the halo
    planes are then copied from the ranks that own them (`exchange_halos`, point-to-point over
    NCCL / NVLink) -- so the next blur, the extrema scan, the orientation and descriptor
    windows of the owned voxels read the same values as on one device;
  * the two global scalars of the path, max|image| (imutil.c:681-695) and max|DoG| per level
    (sift.c:775-783), are max-all-reduced -- max is exact, so the thresholds are identical;
  * octaves K.. are too small to shard: their seed level (K, -1) is gathered from the owners
    and every rank runs them redundantly through sift3d_cuda_detect_from_seed (rank 0's copy
    is the one reported);
  * candidates are merged in the reference's scan order (o, s, z, y, x) = (o, s, rank, local
    order), then compacted like sift.c:1148-1166 (including the strength slot quirk).

The blur forms its z interpolation coordinates (imutil.c:803-853: z - d*unit_factor, the
right-boundary mirror 2*(nz-1) - c - 0.1f) in the coordinates of the whole volume -- the slab
description carries the volume plane of array plane 0 -- so the fractions round exactly as on
a single device and the results are bit-identical for any voxel spacing.

Host-side logic only (layout, exchange, merge): all arithmetic of the path stays in the CUDA
library. Works on CPU tensors with the gloo backend for the layout/exchange/merge tests.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np

DESC_SIG_FCTR = 7.071067812  # sift.c:43
DESC_RAD_FCTR = 2.0          # sift.c:44
ORI_SIG_FCTR = 1.5           # sift.c:41
ORI_RAD_FCTR = 3.0           # sift.c:42


@dataclass(frozen=True)
class Slab:
    """Plane ranges of one rank, in global octave-0 planes."""
    g0: int  # owned [g0, g1)
    g1: int
    a0: int  # held  [a0, a1)
    a1: int

    def own(self, o):
        """Owned planes of octave o in slab-array coordinates."""
        return (self.g0 >> o) - (self.a0 >> o), (self.g1 >> o) - (self.a0 >> o)

    def held(self, o):
        """Held planes of octave o in global octave-o coordinates."""
        return self.a0 >> o, self.a1 >> o

    def owned(self, o):
        return self.g0 >> o, self.g1 >> o


def required_halo(sigma0, num_kp_levels, uz, sharded_octaves, max_half_width=0):
    """Halo planes (octave-0 planes, multiple of 2^K) that cover, at every sharded octave, the
    widest window of the path: the descriptor window radius desc_rad_fctr * desc_sig_fctr * sd
    / unit (sift.c:1453-1461; +1 for the central differences, +1 for ceil), the orientation
    window (sift.c:936), the extrema neighbourhood and the blur's reach."""
    K = sharded_octaves
    sd_rel = sigma0 * 2.0 ** ((num_kp_levels - 1) / num_kp_levels)  # sd / 2^o of the top keypoint level
    w_desc = DESC_RAD_FCTR * DESC_SIG_FCTR * sd_rel / uz
    w_ori = ORI_RAD_FCTR * ORI_SIG_FCTR * sd_rel / uz
    w = int(math.ceil(max(w_desc, w_ori))) + 3
    w = max(w, int(math.ceil(max_half_width / uz)) + 2)
    align = 1 << K
    h = w << (K - 1) if K >= 1 else w
    return (h + align - 1) // align * align


def slab_layout(nz, world, sharded_octaves, halo):
    """Split nz planes into `world` owned ranges on multiples of 2^K, each with its halo."""
    align = 1 << sharded_octaves
    if halo % align:
        raise ValueError("halo must be a multiple of 2^sharded_octaves")
    nblk = nz // align
    if nblk < world:
        raise ValueError(f"{nz} planes cannot be split into {world} slabs aligned to {align}")
    cuts = [align * ((r * nblk) // world) for r in range(world)] + [nz]
    out = []
    for r in range(world):
        g0, g1 = cuts[r], cuts[r + 1]
        out.append(Slab(g0, g1, max(0, g0 - halo), min(nz, g1 + halo)))
    return out


def exchange_plan(layout, o):
    """Point-to-point copies that fill every rank's halo planes of an octave-o level.
    Returns [(src_rank, dst_rank, p0, p1)] in global octave-o planes, deterministic order."""
    plan = []
    for dst, sd in enumerate(layout):
        h0, h1 = sd.held(o)
        for src, ss in enumerate(layout):
            if src == dst:
                continue
            o0, o1 = ss.owned(o)
            p0, p1 = max(h0, o0), min(h1, o1)
            if p0 < p1:
                plan.append((src, dst, p0, p1))
    return plan


def _staged(t, group):
    """CUDA tensors under a backend without device point-to-point (gloo: the single-GPU test
    rig) travel through host copies; under NCCL they go GPU to GPU over NVLink."""
    import torch.distributed as dist

    return t.is_cuda and dist.get_backend(group) != "nccl"


def exchange_halos(level, layout, o, rank, group=None):
    """Fill the non-owned planes of `level` (torch tensor [planes, ny, nx] holding this rank's
    slab of an octave-o level) from the ranks that own them."""
    import torch
    import torch.distributed as dist

    base = layout[rank].held(o)[0]
    staged = _staged(level, group)
    ops, landing = [], []
    for src, dst, p0, p1 in exchange_plan(layout, o):
        view = level[p0 - base:p1 - base] if rank in (src, dst) else None
        if rank == src:
            ops.append(dist.P2POp(dist.isend, view.cpu() if staged else view, dst, group=group))
        elif rank == dst:
            buf = torch.empty(view.shape, dtype=view.dtype) if staged else view
            landing.append((view, buf))
            ops.append(dist.P2POp(dist.irecv, buf, src, group=group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    if staged:
        for view, buf in landing:
            view.copy_(buf)


def gather_owned(level, layout, o, rank, full, group=None):
    """Assemble the whole octave-o level in `full` ([global planes, ny, nx]) on every rank from
    the owned planes of each rank's slab `level`."""
    import torch.distributed as dist

    base = layout[rank].held(o)[0]
    staged = _staged(full, group)
    for src, s in enumerate(layout):
        p0, p1 = s.owned(o)
        if p0 >= p1:
            continue
        if src == rank:
            full[p0:p1].copy_(level[p0 - base:p1 - base])
        if staged:
            buf = full[p0:p1].cpu()
            dist.broadcast(buf, src=src, group=group)
            if src != rank:
                full[p0:p1].copy_(buf)
        else:
            dist.broadcast(full[p0:p1], src=src, group=group)


def reduce_max(vals, group=None):
    """In-place element-wise max over the ranks."""
    import torch.distributed as dist

    if _staged(vals, group):
        buf = vals.cpu()
        dist.all_reduce(buf, op=dist.ReduceOp.MAX, group=group)
        vals.copy_(buf)
    else:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX, group=group)


def merge_candidates(per_rank, tail=None):
    """per_rank[r]: structured array (api.CUDA_KEYPOINT_DTYPE) of rank r's candidates in its
    own scan order, z already global; tail: candidates of the non-sharded octaves. Returns
    all candidates in the reference's scan order (o, s, z, y, x) (sift.c:760-870): within a
    level the ranks' z ranges are disjoint and increasing, so the order is (o, s, rank)."""
    parts = []
    for r, c in enumerate(per_rank):
        if len(c):
            key = np.empty(len(c), dtype=[("o", "i8"), ("s", "i8"), ("r", "i8"), ("i", "i8")])
            key["o"], key["s"], key["r"], key["i"] = c["o"], c["s"], r, np.arange(len(c))
            parts.append((key, c))
    if tail is not None and len(tail):
        key = np.empty(len(tail), dtype=[("o", "i8"), ("s", "i8"), ("r", "i8"), ("i", "i8")])
        key["o"], key["s"], key["r"], key["i"] = tail["o"], tail["s"], 0, np.arange(len(tail))
        parts.append((key, tail))
    if not parts:
        return np.zeros(0, dtype=per_rank[0].dtype if per_rank else None), np.zeros(0, np.int64)
    keys = np.concatenate([k for k, _ in parts])
    cands = np.concatenate([c for _, c in parts])
    order = np.argsort(keys, order=("o", "s", "r", "i"), kind="stable")
    return cands[order], order


def compact_keypoints(cands):
    """assign_orientations' in-order compaction (sift.c:1148-1166): survivors keep candidate
    order; copy_Keypoint does not copy `strength` (sift.c:372-384), so survivor j carries the
    strength of candidate j."""
    keep = np.flatnonzero(cands["valid"] != 0)
    out = cands[keep].copy()
    out["strength"] = cands["strength"][:len(keep)]
    return out, keep


class _DevArray:
    """Minimal __cuda_array_interface__ carrier for a raw float32 device pointer."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(int(v) for v in shape), "typestr": "<f4",
                                         "data": (int(ptr), False), "version": 2, "strides": None}


def _wrap(ptr, shape, device):
    import torch

    return torch.as_tensor(_DevArray(ptr, shape), device=torch.device("cuda", device))


class SlabDetector:
    """detect + describe of one (nx, ny, nz) volume split in z over the ranks of a
    torch.distributed process group (one process per GPU, backend "nccl").

    Parameters mirror the reference's detector setters (sift.c:498-561). `sharded_octaves`
    (K) octaves run on slabs; the remaining octaves run replicated on every rank."""

    def __init__(self, dims, units=(1.0, 1.0, 1.0), sharded_octaves=2, halo=None, device=None,
                 group=None, peak_thresh=None, corner_thresh=None, num_kp_levels=None,
                 sigma_n=None, sigma0=None):
        import torch
        import torch.distributed as dist
        from . import api

        self._api = api
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.nx, self.ny, self.nz = (int(v) for v in dims)
        self.units = tuple(float(u) for u in units)
        self.peak_thresh = 0.1 if peak_thresh is None else float(peak_thresh)      # sift.c:36
        self.corner_thresh = 0.4 if corner_thresh is None else float(corner_thresh)  # sift.c:38
        self.num_kp_levels = 3 if num_kp_levels is None else int(num_kp_levels)   # sift.c:37
        self.sigma_n = 1.15 if sigma_n is None else float(sigma_n)                # sift.c:39
        self.sigma0 = 1.6 if sigma0 is None else float(sigma0)                    # sift.c:40

        # parameter holder: the host C library designs scales and filters (fill_plan)
        self._det = api.Detector(self.device)
        for fn, v in ((self._det.set_peak_thresh, self.peak_thresh),
                      (self._det.set_corner_thresh, self.corner_thresh),
                      (self._det.set_num_kp_levels, self.num_kp_levels),
                      (self._det.set_sigma_n, self.sigma_n), (self._det.set_sigma0, self.sigma0)):
            if fn(v):
                raise api.Sift3DError("invalid detector parameter")
        self._L, self._C = api.lib(), api.libcuda()

        total = int(math.log2(min(self.nx, self.ny, self.nz))) - 3 + 1  # sift.c:441-452
        if total < 1:
            raise api.Sift3DError("image too small: at least 8 voxels in each dimension")
        self.total_octaves = total
        self.K = K = max(1, min(int(sharded_octaves), total))
        self.tail_octaves = total - K

        plan = api.CudaPlanDesc()
        # filters do not depend on the geometry: design once to learn the blur's reach
        if self._L.sift3d_detector_fill_plan(self._det.handle, self.nx, self.ny, self.nz,
                                             *self.units, 0, 1, C.byref(plan)):
            raise api.Sift3DError("fill_plan failed")
        hw = max([plan.first.width // 2] + [plan.octave[i].width // 2
                                            for i in range(self.num_kp_levels + 2)])
        need = required_halo(self.sigma0, self.num_kp_levels, self.units[2], K, hw)
        self.halo = need if halo is None else int(halo)
        if self.halo < need:
            raise ValueError(f"halo {self.halo} is smaller than the path's widest window ({need})")
        self.layout = slab_layout(self.nz, self.world, K, self.halo)
        self.slab = self.layout[self.rank]
        self.slab_nz = self.slab.a1 - self.slab.a0

        # slab context: K sharded octaves (+ one more whose level -1 receives the seed)
        self.ctx = self._C.sift3d_cuda_ctx_create(self.device)
        if not self.ctx:
            raise api.Sift3DError("no usable CUDA device; there is no CPU path")
        noct_plan = K + (1 if self.tail_octaves else 0)
        if self._L.sift3d_detector_fill_plan(self._det.handle, self.nx, self.ny, self.slab_nz,
                                             *self.units, 0, noct_plan, C.byref(plan)) or \
                self._C.sift3d_cuda_plan(self.ctx, C.byref(plan)):
            raise api.Sift3DError("planning the slab context failed")
        self.slab_desc = api.CudaSlabDesc()
        self.slab_desc.num_sharded_octaves = K
        for o in range(noct_plan):
            self.slab_desc.own0[o], self.slab_desc.own1[o] = self.slab.own(o)
            self.slab_desc.org[o] = self.slab.held(o)[0]

        # replicated context for the octaves that are too small to shard
        self.tail_ctx = None
        if self.tail_octaves:
            self.tail_dims = (self.nx >> K, self.ny >> K, self.nz >> K)
            self.tail_units = tuple(math.ldexp(u, K) for u in self.units)
            self.tail_ctx = self._C.sift3d_cuda_ctx_create(self.device)
            if self._L.sift3d_detector_fill_plan(self._det.handle, *self.tail_dims,
                                                 *self.tail_units, K, self.tail_octaves,
                                                 C.byref(plan)) or \
                    self._C.sift3d_cuda_plan(self.tail_ctx, C.byref(plan)) or \
                    self._C.sift3d_cuda_set_base_octave(self.tail_ctx, K):
                raise api.Sift3DError("planning the tail context failed")
            self._seed = torch.empty((self.tail_dims[2], self.tail_dims[1], self.tail_dims[0]),
                                     dtype=torch.float32, device=torch.device("cuda", self.device))

        self._error = None
        self.exchanged_bytes = 0
        self._cb = api.CudaSlabCallbacks(None, api.REDUCE_MAX_FN(self._reduce_max),
                                         api.EXCHANGE_FN(self._exchange))

    # ---- callbacks from sift3d_cuda_detect_slab (context stream drained) ----
    def _reduce_max(self, user, d_vals, count):
        import torch
        import torch.distributed as dist

        try:
            t = _wrap(d_vals, (count,), self.device)
            reduce_max(t, self.group)
            torch.cuda.current_stream(self.device).synchronize()
            return 0
        except Exception as e:  # noqa: BLE001 - reported through the C status
            self._error = e
            return -1

    def _exchange(self, user, o, s, d_level, nx, ny, nz, own0, own1):
        import torch

        try:
            if o >= self.K:
                return 0  # the seed of the replicated octaves is gathered whole, below
            t = _wrap(d_level, (nz, ny, nx), self.device)
            exchange_halos(t, self.layout, o, self.rank, self.group)
            torch.cuda.current_stream(self.device).synchronize()
            self.exchanged_bytes += 4 * nx * ny * (nz - (own1 - own0))
            return 0
        except Exception as e:  # noqa: BLE001
            self._error = e
            return -1

    # ---- the path ----
    def _fetch(self, ctx, n):
        buf = np.zeros(max(n, 1), dtype=self._api.CUDA_KEYPOINT_DTYPE)
        got = self._C.sift3d_cuda_fetch_candidates(ctx, buf.ctypes.data_as(C.c_void_p), n)
        if got != n:
            raise self._api.Sift3DError("sift3d_cuda_fetch_candidates failed")
        return buf[:n]

    def _describe(self, ctx, cands):
        """768-bin descriptors of the valid candidates; one row per candidate (zeros where
        the candidate did not survive orientation assignment)."""
        out = np.zeros((len(cands), 768), dtype=np.float32)
        keep = np.flatnonzero(cands["valid"] != 0)
        if len(keep):
            keys = np.ascontiguousarray(cands[keep])
            desc = np.empty((len(keep), 768), dtype=np.float32)
            if self._C.sift3d_cuda_describe(ctx, keys.ctypes.data_as(C.c_void_p), len(keep),
                                            desc.ctypes.data_as(C.c_void_p), 0):
                raise self._api.Sift3DError("sift3d_cuda_describe failed")
            out[keep] = desc
        return out

    def detect_and_describe(self, slab_image):
        """slab_image: this rank's held planes [a0, a1) of the volume, float32 [a1-a0, ny, nx]
        (numpy array on the host, or a torch CUDA tensor on this rank's device).
        Returns on rank 0 a dict(candidates, keypoints, descriptors); None on other ranks."""
        import torch
        import torch.distributed as dist

        api = self._api
        on_device = isinstance(slab_image, torch.Tensor)
        if on_device:
            if not (slab_image.is_cuda and slab_image.is_contiguous() and
                    slab_image.dtype == torch.float32):
                raise ValueError("device slab must be a contiguous float32 CUDA tensor")
            torch.cuda.current_stream(self.device).synchronize()
            ptr = slab_image.data_ptr()
        else:
            slab_image = np.ascontiguousarray(slab_image, dtype=np.float32)
            ptr = slab_image.ctypes.data
        if tuple(slab_image.shape) != (self.slab_nz, self.ny, self.nx):
            raise ValueError(f"slab must be {(self.slab_nz, self.ny, self.nx)}, got "
                             f"{tuple(slab_image.shape)}")
        units = (C.c_double * 3)(*self.units)
        n = C.c_int(0)
        self._error = None
        rc = self._C.sift3d_cuda_detect_slab(self.ctx, ptr, int(on_device), units, self.peak_thresh,
                                             self.corner_thresh, C.byref(self.slab_desc),
                                             C.byref(self._cb), C.byref(n))
        if self._error is not None:
            raise self._error
        if rc:
            raise api.Sift3DError("sift3d_cuda_detect_slab failed")
        local = self._fetch(self.ctx, n.value)
        desc = self._describe(self.ctx, local)
        cands = local.copy()
        for o in range(self.K):  # slab-array z -> volume z, in octave coordinates
            cands["z"][cands["o"] == o] += self.slab.held(o)[0]

        tail = tail_desc = None
        if self.tail_ctx is not None:
            K = self.K
            d = self._C.sift3d_cuda_level_ptr(self.ctx, K, -1)
            lvl = _wrap(d, (self.slab_nz >> K, self.ny >> K, self.nx >> K), self.device)
            gather_owned(lvl, self.layout, K, self.rank, self._seed, self.group)
            torch.cuda.current_stream(self.device).synchronize()
            tunits = (C.c_double * 3)(*self.tail_units)
            if self._C.sift3d_cuda_detect_from_seed(self.tail_ctx, self._seed.data_ptr(), 1, tunits,
                                                    self.peak_thresh, self.corner_thresh,
                                                    C.byref(n)):
                raise api.Sift3DError("sift3d_cuda_detect_from_seed failed")
            if self.rank == 0:
                tail = self._fetch(self.tail_ctx, n.value)
                tail_desc = self._describe(self.tail_ctx, tail)

        gathered = [None] * self.world
        dist.all_gather_object(gathered, (cands, desc), group=self.group)
        if self.rank != 0:
            return None
        per_rank = [g[0] for g in gathered]
        rows = [g[1] for g in gathered if len(g[0])]
        if tail is not None and len(tail):
            rows.append(tail_desc)
        merged, order = merge_candidates(per_rank, tail)
        all_desc = np.concatenate(rows)[order] if rows else np.zeros((0, 768), np.float32)
        keypoints, keep = compact_keypoints(merged)
        return {"candidates": merged, "keypoints": keypoints, "descriptors": all_desc[keep]}

    def close(self):
        for name in ("ctx", "tail_ctx"):
            h = getattr(self, name, None)
            if h:
                self._C.sift3d_cuda_ctx_destroy(h)
                setattr(self, name, None)

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass
