"""Multi-GPU host: one process per GPU, samples sharded, grasps all-gathered.

The path shards by grasp sample (samples are independent in every stage after the
normals; reference hand_search.cpp:118,194, learning.cpp:12).  Per detect call:

  1. rank 0 holds the cloud; xyz (12 B/pt) + camera mask (1 B/pt) + the sample list are
     broadcast with torch.distributed (NCCL over NVLink on GPUs);
  2. every rank builds the identical cell index (cheaper than shipping it) and evaluates the
     STRIDED shard rank, rank + world, ... of the sample list (sample lists are sorted by point
     index, i.e. spatially coherent: contiguous blocks would give the ranks different parts of
     the scene and unequal work); `set_sample_id_map(rank, world)` keeps the counter-based draws
     identical to a single-GPU run of the whole list;
  3. per-rank antipodal lists (192-byte records) are all-gathered (counts first, then
     records padded to the maximum count) and the global top-`num_selected` is taken with the
     same ordering rule as the single-GPU path (score descending, list order on ties).  When a
     top-k is requested the global top-k lies inside the union of the per-rank top-k lists, so
     each rank contributes its SELECTED list in one fixed-size all-gather.

On GPUs the whole exchange stays in device memory: ONE broadcast of a packed input buffer
(`pack_inputs`: xyz | camera mask | samples in one allocation), the sample block handed to the
library by device pointer, the selected records copied device-to-device into the all-gather send
buffer, one device->host read of the gathered records at the end.

There is no other data-path collective.  The same code runs over `gloo` with CPU tensors for
the world_size-2 CPU tests (tests drive it with the oracle context; the product uses libag2).
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


def pack_inputs(xyz, cam, samples, device):
    """One device allocation holding xyz [N,3] f32 | samples [S] i32 | cam [N] u8 (4-byte aligned
    sections); returns (packed u8 tensor, xyz view, cam view, samples view).  Rank 0 keeps its
    resident cloud in this form so that a detect call needs a single broadcast."""
    n, s = int(xyz.shape[0]), int(samples.shape[0])
    packed = torch.empty(12 * n + 4 * s + n, dtype=torch.uint8, device=device)
    xv = packed[: 12 * n].view(torch.float32).view(n, 3)
    sv = packed[12 * n: 12 * n + 4 * s].view(torch.int32)
    cv = packed[12 * n + 4 * s:]
    if xyz is not None:
        xv.copy_(torch.as_tensor(xyz).reshape(n, 3), non_blocking=True)
        sv.copy_(torch.as_tensor(samples), non_blocking=True)
        cv.copy_(torch.as_tensor(cam), non_blocking=True)
    return packed, xv, cv, sv


def _packed_views(packed, n, s):
    return (packed[: 12 * n].view(torch.float32).view(n, 3), packed[12 * n + 4 * s:],
            packed[12 * n: 12 * n + 4 * s].view(torch.int32))


class ShardedDetector:
    def __init__(self, ctx: capi.Context, rank: int = 0, world: int = 1, device=None, group=None,
                 shard_normals: bool = True):
        self.ctx, self.rank, self.world, self.group = ctx, rank, world, group
        self.shard_normals = shard_normals  # GPU fast path: normals computed 1/world per rank + all-gather
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

    def detect_device(self, xyz_d, cam_d, num_cams, samples_d, shape_hint=None):
        """Cloud resident in HBM on rank 0 (torch CUDA tensors)."""
        xyz_d, cam_d, num_cams, samples_d = self._broadcast_inputs(xyz_d, cam_d, num_cams, samples_d, shape_hint)
        if self.world > 1:
            torch.cuda.current_stream().synchronize()
        self.ctx.set_cloud_device(xyz_d.data_ptr(), xyz_d.shape[0], cam_d.data_ptr(), num_cams)
        return self._run_block(samples_d.cpu().numpy())

    def detect_packed(self, packed_d, n, s, num_cams):
        """GPU fast path.  packed_d: the `pack_inputs` buffer (valid on rank 0, a same-sized receive
        buffer elsewhere); (n, s, num_cams) known on every rank.  One broadcast, device-resident
        sample block, device-side all-gather of the per-rank top-k."""
        import time
        p = self.ctx.params
        topk = p.num_selected > 0 and p.antipodal_mode != capi.MODE_NONE
        t0 = time.perf_counter()
        if self.world > 1:
            dist.broadcast(packed_d, src=0, group=self.group)
            torch.cuda.current_stream().synchronize()
        t1 = time.perf_counter()
        xyz_d, cam_d, smp_d = _packed_views(packed_d, n, s)
        if not topk:  # unbounded lists: the general path
            self.ctx.set_cloud_device(xyz_d.data_ptr(), n, cam_d.data_ptr(), num_cams)
            return self._run_block(smp_d.cpu().numpy())
        self.ctx.set_cloud_device(xyz_d.data_ptr(), n, cam_d.data_ptr(), num_cams)
        if self.world > 1 and self.shard_normals:
            # the normals of the whole cloud are needed by every rank: each computes 1/world of them
            # (interleaved 32-point granules of the cell-sorted cloud), one all-gather installs the rest
            cap = self.ctx.normals_shard_points(self.world)
            if getattr(self, "_nrm_cap", None) != cap:
                self._nrm_tx = torch.empty(cap * 4, dtype=torch.float32, device=self.device)
                self._nrm_rx = torch.empty(self.world * cap * 4, dtype=torch.float32, device=self.device)
                self._nrm_cap = cap
            self.ctx.compute_normals_shard(self.rank, self.world, self._nrm_tx.data_ptr(), cap)
            dist.all_gather_into_tensor(self._nrm_rx, self._nrm_tx, group=self.group)
            torch.cuda.current_stream().synchronize()
            self.ctx.set_normals_shards(self._nrm_rx.data_ptr(), cap, self.world)
        t1b = time.perf_counter()
        mine = smp_d[self.rank::self.world].contiguous() if self.world > 1 else smp_d
        n_mine = shard_size(s, self.world, self.rank)
        if self.world > 1:
            torch.cuda.current_stream().synchronize()
        self.ctx.set_sample_indices_device(mine.data_ptr(), n_mine)
        self.ctx.set_sample_id_map(self.rank, self.world)
        rec, cap = capi.HYPOTHESIS_DTYPE.itemsize, p.num_selected
        if getattr(self, "_tx_cap", None) != cap:
            self._tx = torch.empty(cap * rec, dtype=torch.uint8, device=self.device)
            self._rxg = torch.empty(self.world * cap * rec, dtype=torch.uint8, device=self.device)
            self._tx_cap = cap
        if self.world == 1:  # one rank: the device ranking already is the final order, read it directly
            self.ctx.detect()
            out = self.ctx.hypotheses(capi.LIST_SELECTED)
            t2 = time.perf_counter()
            self.last_ms = {"bcast": (t1 - t0) * 1e3, "normals": (t1b - t1) * 1e3, "detect": (t2 - t1b) * 1e3,
                            "gather": 0.0}
            return out
        if n_mine:
            self.ctx.detect()
            self.ctx.copy_hypotheses_device(capi.LIST_SELECTED, self._tx.data_ptr(), cap)
        else:
            self._tx.zero_()
            torch.cuda.current_stream().synchronize()
        t2 = time.perf_counter()
        dist.all_gather_into_tensor(self._rxg, self._tx, group=self.group)
        allr = self._rxg.cpu().numpy().view(capi.HYPOTHESIS_DTYPE).reshape(self.world, cap)
        t3 = time.perf_counter()
        # host-side wall time of the three segments of the last call (ms): input exchange, local
        # detect (includes waiting for nothing but this rank's GPU), result exchange
        self.last_ms = {"bcast": (t1 - t0) * 1e3, "normals": (t1b - t1) * 1e3, "detect": (t2 - t1b) * 1e3,
                        "gather": (t3 - t2) * 1e3}
        return capi.merge_selected_blocks(allr, p.num_selected)  # native: the same rule as merge_selected

    def detect_host(self, xyz_h, cam_h, num_cams, samples_h, shape_hint=None):
        """Host buffers (pinned torch CPU tensors on rank 0): H2D inside the call."""
        if self.world == 1:
            self.ctx.set_cloud(xyz_h.numpy(), cam_h.numpy(), num_cams)
            return self._run_block(samples_h.numpy())
        if shape_hint is not None:  # sizes known everywhere: stage into the packed buffer, one broadcast
            n, s, num_cams = shape_hint
            if getattr(self, "_pk_key", None) != (n, s):
                self._pk = torch.empty(12 * n + 4 * s + n, dtype=torch.uint8, device=self.device)
                self._pk_key = (n, s)
            if self.rank == 0:
                xv, cv, sv = _packed_views(self._pk, n, s)
                xv.copy_(xyz_h, non_blocking=True)
                cv.copy_(cam_h, non_blocking=True)
                sv.copy_(samples_h, non_blocking=True)
            return self.detect_packed(self._pk, n, s, num_cams)
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
