"""Multi-GPU host: one process per GPU, samples sharded, grasps all-gathered.

The path shards by grasp sample (samples are independent in every stage after the
normals; reference hand_search.cpp:118,194, learning.cpp:12).  Per detect call:

  1. rank 0 holds the cloud; xyz (12 B/pt) + camera mask (1 B/pt) + the sample list are
     broadcast with torch.distributed (NCCL over NVLink on GPUs);
  2. every rank builds the identical cell index (cheaper than shipping it) and evaluates the
     contiguous block [lo, hi) of the sample list; `set_sample_id_offset(lo)` keeps the
     counter-based draws identical to a single-GPU run of the whole list;
  3. per-rank antipodal lists (192-byte records) are all-gathered (counts first, then
     records padded to the maximum count) and the global top-`num_selected` is taken with the
     same ordering rule as the single-GPU path (score descending, list order on ties).

There is no other data-path collective.  The same code runs over `gloo` with CPU tensors for
the world_size-2 CPU tests (tests drive it with the oracle context; the product uses libag2).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import capi


def partition(n: int, world: int, rank: int):
    """Contiguous block of rank `rank` when n samples are split over `world` ranks."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def merge_selected(parts, num_selected: int, keep_all: bool = False) -> np.ndarray:
    """Concatenates per-rank antipodal lists (rank order == global sample order) and takes the
    top `num_selected` by (score desc, list position asc) — grasp_detector.cpp:238-252."""
    allh = np.concatenate(parts) if parts else np.zeros(0, capi.HYPOTHESIS_DTYPE)
    if keep_all or len(allh) == 0:
        return allh
    order = np.argsort(-allh["score"], kind="stable")
    if num_selected > 0:
        order = order[:num_selected]
    return allh[order]


class ShardedDetector:
    def __init__(self, ctx: capi.Context, rank: int = 0, world: int = 1, device=None, group=None):
        self.ctx, self.rank, self.world, self.group = ctx, rank, world, group
        self.device = device if device is not None else torch.device("cpu")

    # ---- collectives -----------------------------------------------------------------
    def _bcast(self, t):
        if self.world > 1:
            dist.broadcast(t, src=0, group=self.group)
        return t

    def _broadcast_inputs(self, xyz, cam, num_cams, samples, shape_hint=None):
        """xyz [N,3] f32, cam [N] u8, samples [S] i32 tensors on self.device on rank 0.
        shape_hint = (N, S, num_cams) known on every rank skips the size broadcast (and its
        device->host sync); receive buffers are kept between calls."""
        if self.world == 1:
            return xyz, cam, num_cams, samples
        if shape_hint is None:
            meta = torch.zeros(3, dtype=torch.int64, device=self.device)
            if self.rank == 0:
                meta[0], meta[1], meta[2] = xyz.shape[0], samples.shape[0], num_cams
            self._bcast(meta)
            n, s, num_cams = int(meta[0]), int(meta[1]), int(meta[2])
        else:
            n, s, num_cams = shape_hint
        if self.rank != 0:
            key = (n, s)
            if getattr(self, "_rx_key", None) != key:
                self._rx = (torch.empty((n, 3), dtype=torch.float32, device=self.device),
                            torch.empty((n,), dtype=torch.uint8, device=self.device),
                            torch.empty((s,), dtype=torch.int32, device=self.device))
                self._rx_key = key
            xyz, cam, samples = self._rx
        self._bcast(xyz)
        self._bcast(cam)
        self._bcast(samples)
        return xyz, cam, num_cams, samples

    def _gather_fixed(self, local: np.ndarray, cap: int):
        """One all-gather of fixed-size blocks: [count | cap records] per rank (used when every
        rank contributes at most `cap` = num_selected records)."""
        rec = capi.HYPOTHESIS_DTYPE.itemsize
        buf = np.zeros(16 + cap * rec, dtype=np.uint8)
        buf[:8] = np.frombuffer(np.int64(len(local)).tobytes(), dtype=np.uint8)
        buf[16:16 + len(local) * rec] = local.view(np.uint8).reshape(-1)
        t = torch.from_numpy(buf).to(self.device)
        out = torch.empty(self.world * buf.size, dtype=torch.uint8, device=self.device)
        dist.all_gather_into_tensor(out, t, group=self.group)
        allb = out.cpu().numpy().reshape(self.world, buf.size)
        parts = []
        for r in range(self.world):
            c = int(np.frombuffer(allb[r, :8].tobytes(), dtype=np.int64)[0])
            parts.append(allb[r, 16:16 + c * rec].copy().view(capi.HYPOTHESIS_DTYPE))
        return parts

    def _gather(self, local: np.ndarray):
        if self.world == 1:
            return [local]
        cnt = torch.tensor([len(local)], dtype=torch.int64, device=self.device)
        counts = [torch.zeros_like(cnt) for _ in range(self.world)]
        dist.all_gather(counts, cnt, group=self.group)
        counts = [int(c) for c in counts]
        mx = max(max(counts), 1)
        buf = np.zeros(mx, dtype=capi.HYPOTHESIS_DTYPE)
        buf[: len(local)] = local
        t = torch.from_numpy(buf.view(np.uint8).reshape(-1)).to(self.device)
        outs = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(outs, t, group=self.group)
        return [o.cpu().numpy().view(capi.HYPOTHESIS_DTYPE)[:c] for o, c in zip(outs, counts)]

    # ---- detect ----------------------------------------------------------------------
    def _run_block(self, samples_all: np.ndarray):
        lo, hi = partition(len(samples_all), self.world, self.rank)
        self.lo = lo
        self.ctx.set_sample_indices(samples_all[lo:hi])
        self.ctx.set_sample_id_offset(lo)
        local = np.zeros(0, capi.HYPOTHESIS_DTYPE)
        p = self.ctx.params
        # the global top-k is inside the union of the per-rank top-k, so when a top-k is asked for
        # each rank ships its own SELECTED list (<= num_selected records, one fixed-size all-gather)
        topk = p.num_selected > 0 and p.antipodal_mode != capi.MODE_NONE
        if hi > lo:
            self.ctx.detect()
            which = capi.LIST_SELECTED if (self.world == 1 or topk) else capi.LIST_ANTIPODAL
            local = self.ctx.hypotheses(which).copy()
            local["sample_id"] += lo
        if self.world == 1:
            return local
        if topk:
            return merge_selected(self._gather_fixed(local, p.num_selected), p.num_selected)
        return merge_selected(self._gather(local), p.num_selected, keep_all=p.antipodal_mode == capi.MODE_NONE)

    def detect_device(self, xyz_d, cam_d, num_cams, samples_d, shape_hint=None):
        """Cloud resident in HBM on rank 0 (torch CUDA tensors)."""
        xyz_d, cam_d, num_cams, samples_d = self._broadcast_inputs(xyz_d, cam_d, num_cams, samples_d, shape_hint)
        if self.world > 1:
            torch.cuda.current_stream().synchronize()
        self.ctx.set_cloud_device(xyz_d.data_ptr(), xyz_d.shape[0], cam_d.data_ptr(), num_cams)
        return self._run_block(samples_d.cpu().numpy())

    def detect_host(self, xyz_h, cam_h, num_cams, samples_h, shape_hint=None):
        """Host buffers (pinned torch CPU tensors on rank 0): H2D inside the call."""
        if self.world == 1:
            self.ctx.set_cloud(xyz_h.numpy(), cam_h.numpy(), num_cams)
            return self._run_block(samples_h.numpy())
        if self.rank == 0:
            xyz_d, cam_d, smp_d = (xyz_h.to(self.device, non_blocking=True), cam_h.to(self.device, non_blocking=True),
                                   samples_h.to(self.device, non_blocking=True))
        else:
            xyz_d = cam_d = smp_d = None
        return self.detect_device(xyz_d, cam_d, num_cams, smp_d, shape_hint)

    def detect_numpy(self, xyz, cam, num_cams, samples):
        """Backend-neutral entry (gloo CPU tests): numpy inputs on rank 0, None elsewhere."""
        if self.rank == 0:
            tx, tc, ts = (torch.from_numpy(np.ascontiguousarray(xyz)), torch.from_numpy(np.ascontiguousarray(cam)),
                          torch.from_numpy(np.ascontiguousarray(samples, dtype=np.int32)))
        else:
            tx = tc = ts = None
        tx, tc, num_cams, ts = self._broadcast_inputs(tx, tc, num_cams, ts)
        self.ctx.set_cloud(tx.numpy(), tc.numpy(), num_cams)
        return self._run_block(ts.numpy())
