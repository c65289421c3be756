"""Host logic of the sample sharding, restated in Python for the CPU tests.

The multi-GPU data path is native: csrc/group.cu behind ag2_group_* (include/ag2.h) owns the NCCL
communicator, the peer-memory windows and the exchanges; `capi.Group` binds it and bench.py /
tests/multi_gpu_check.py call that.  This module keeps the SAME host logic -- strided sample shards
(rank, rank + world, ...), `set_sample_id_map(rank, world)` so that a shard draws like the whole list,
per-rank top-k lists merged with the single-GPU ordering rule (grasp_detector.cpp:238-252), importance-sampling
rounds in which every rank draws the same batch from the merged hands -- over `torch.distributed` with CPU
tensors, so that the world_size-2 `gloo` tests can run it against the oracle context on a box without GPUs
(tests/test_sharded_gloo.py).  It is not on the product's GPU path.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import capi


def partition(n: int, world: int, rank: int):
    """Contiguous block of rank `rank` when n items are split over `world` ranks."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_size(n: int, world: int, rank: int) -> int:
    """Number of samples in the strided shard rank, rank + world, ... of n samples."""
    return (n - rank + world - 1) // world if n > rank else 0


def merge_selected(parts, num_selected: int, keep_all: bool = False) -> np.ndarray:
    """Merges per-rank lists (sample_id already global) back into the single-process list order
    (sample, orientation) and takes the top `num_selected` by (score desc, list position asc) —
    grasp_detector.cpp:238-252."""
    allh = np.concatenate(parts) if parts else np.zeros(0, capi.HYPOTHESIS_DTYPE)
    if len(allh):
        allh = allh[np.lexsort((allh["orientation_id"], allh["sample_id"]))]
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
        mine = np.ascontiguousarray(samples_all[self.rank::self.world])
        self.ctx.set_sample_indices(mine)
        self.ctx.set_sample_id_map(self.rank, self.world)
        local = np.zeros(0, capi.HYPOTHESIS_DTYPE)
        p = self.ctx.params
        # the global top-k is inside the union of the per-rank top-k, so when a top-k is asked for
        # each rank ships its own SELECTED list (<= num_selected records, one fixed-size all-gather)
        topk = p.num_selected > 0 and p.antipodal_mode != capi.MODE_NONE
        if len(mine):
            self.ctx.detect()
            which = capi.LIST_SELECTED if (self.world == 1 or topk) else capi.LIST_ANTIPODAL
            local = self.ctx.hypotheses(which).copy()
            local["sample_id"] = self.rank + self.world * local["sample_id"]
        if self.world == 1:
            return local
        if topk:
            return merge_selected(self._gather_fixed(local, p.num_selected), p.num_selected)
        return merge_selected(self._gather(local), p.num_selected, keep_all=p.antipodal_mode == capi.MODE_NONE)

    def _run_block_points(self, points_all: np.ndarray):
        """Same as _run_block for xyz samples (importance-sampling rounds)."""
        mine = np.ascontiguousarray(points_all[self.rank::self.world])
        p = self.ctx.params
        topk = p.num_selected > 0 and p.antipodal_mode != capi.MODE_NONE
        local = np.zeros(0, capi.HYPOTHESIS_DTYPE)
        if len(mine):
            self.ctx.set_sample_points(mine)
            self.ctx.set_sample_id_map(self.rank, self.world)
            self.ctx.detect()
            which = capi.LIST_SELECTED if (self.world == 1 or topk) else capi.LIST_ANTIPODAL
            local = self.ctx.hypotheses(which).copy()
            local["sample_id"] = self.rank + self.world * local["sample_id"]
        if self.world == 1:
            return local
        if topk:
            return merge_selected(self._gather_fixed(local, p.num_selected), p.num_selected)
        return merge_selected(self._gather(local), p.num_selected, keep_all=p.antipodal_mode == capi.MODE_NONE)

    def detect_importance_numpy(self, xyz, cam, num_cams, samples, ip):
        """ImportanceSampling::detectGraspPoses (importance_sampling.cpp:30-103) over all ranks: the
        cloud is broadcast once and its index and normals stay resident on every rank; every round's
        samples are sharded like a plain detect call and the per-round selections merged, so all
        ranks hold the same hands found so far and draw the SAME next batch from them (counter-based
        stream; no collective for the sampler).  Returns the hands of all rounds (the single-GPU
        LIST_IMPORTANCE: per-round selection order, sample_id counting through the rounds).  The
        final clustering of the reference (:105-109) is left to the caller."""
        if self.ctx.params.min_inliers > 0:
            raise ValueError("sharded importance sampling returns the unclustered hands: set min_inliers = 0")
        if self.rank == 0:
            tx, tc, ts = (torch.from_numpy(np.ascontiguousarray(xyz)).to(self.device),
                          torch.from_numpy(np.ascontiguousarray(cam)).to(self.device),
                          torch.from_numpy(np.ascontiguousarray(samples, dtype=np.int32)).to(self.device))
        else:
            tx = tc = ts = None
        tx, tc, num_cams, ts = self._broadcast_inputs(tx, tc, num_cams, ts)
        self.ctx.set_cloud(tx.cpu().numpy(), tc.cpu().numpy(), num_cams)
        smp = ts.cpu().numpy()
        hands = self._run_block(smp)
        offset = len(smp)
        if len(hands) == 0:
            return hands
        for it in range(int(ip.num_iterations)):
            pts = self.ctx.draw_importance_samples(ip, it, hands)
            sel = self._run_block_points(pts)
            sel["sample_id"] += offset
            offset += len(pts)
            hands = np.concatenate([hands, sel])
        return hands

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
