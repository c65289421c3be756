"""world_size-2 gloo test of the multi-GPU host logic (partition, broadcast, all-gather, merge),
driven with the CPU oracle as the per-rank backend (the product backend needs a GPU)."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

from agile_grasp2_b200 import capi, scenes, sharded
from common import make_params
from oracle import binding as ob

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_covers_everything():
    for n in (0, 1, 7, 8, 1000):
        for w in (1, 2, 3, 8):
            blocks = [sharded.partition(n, w, r) for r in range(w)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(w - 1))
            assert max(b[1] - b[0] for b in blocks) - min(b[1] - b[0] for b in blocks) <= 1
            assert [sharded.shard_size(n, w, r) for r in range(w)] == [len(range(r, n, w)) for r in range(w)]


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = make_params(capi.MODE_GEOMETRIC, num_selected=10, num_threads=2)
    ctx = ob.context(p)
    det = sharded.ShardedDetector(ctx, rank=rank, world=world)
    sc = scenes.make_scene("tiny", num_samples=40) if rank == 0 else None
    sel = det.detect_numpy(*((sc["xyz"], sc["cam_mask"], sc["num_cams"], sc["samples"]) if rank == 0 else (None, None, 0, None)))
    np.save(os.path.join(out_dir, "sel%d.npy" % rank), sel)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])  # 40 samples over three ranks: shards of 14, 13, 13
def test_two_rank_result_equals_single_process(tmp_path, world):
    port = 29541 + (os.getpid() % 200) + 211 * world
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    a = np.load(tmp_path / "sel0.npy")
    for r in range(1, world):
        assert a.tobytes() == np.load(tmp_path / ("sel%d.npy" % r)).tobytes()  # every rank holds the same merged selection
    p = make_params(capi.MODE_GEOMETRIC, num_selected=10, num_threads=2)
    sc = scenes.make_scene("tiny", num_samples=40)
    o = ob.context(p)
    o.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); o.set_sample_indices(sc["samples"]); o.detect()
    ref = o.hypotheses(capi.LIST_SELECTED)
    assert len(a) == len(ref)
    for f in ("sample_id", "orientation_id", "full_antipodal", "finger_slot", "depth_step", "num_in_box"):
        assert np.array_equal(a[f], ref[f]), f
    assert np.array_equal(a["bottom"], ref["bottom"])  # the id map keeps the rand() replacement stream identical


def _is_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = make_params(capi.MODE_NONE, num_selected=10, num_threads=2)
    ctx = ob.context(p)
    ip = ctx.default_is_params()
    ip.num_iterations, ip.num_samples, ip.seed = 2, 30, 4
    det = sharded.ShardedDetector(ctx, rank=rank, world=world)
    sc = scenes.make_scene("tiny", num_samples=40) if rank == 0 else None
    res = det.detect_importance_numpy(*((sc["xyz"], sc["cam_mask"], sc["num_cams"], sc["samples"]) if rank == 0
                                        else (None, None, 0, None)), ip)
    np.save(os.path.join(out_dir, "is%d.npy" % rank), res)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_importance_sampling_equals_single_process(tmp_path):
    """Importance-sampling rounds sharded over two ranks: every rank draws the same batches from the merged
    hands and the accumulated list equals the single-process ag2o_detect_importance list."""
    world, port = 2, 29741 + (os.getpid() % 200)
    mp.spawn(_is_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    a, b = np.load(tmp_path / "is0.npy"), np.load(tmp_path / "is1.npy")
    assert a.tobytes() == b.tobytes()
    p = make_params(capi.MODE_NONE, num_selected=10, num_threads=2)
    sc = scenes.make_scene("tiny", num_samples=40)
    o = ob.context(p)
    ip = o.default_is_params()
    ip.num_iterations, ip.num_samples, ip.seed = 2, 30, 4
    o.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); o.set_sample_indices(sc["samples"])
    o.detect_importance(ip)
    ref = o.hypotheses(capi.LIST_IMPORTANCE)
    assert len(ref) > 10 and a.tobytes() == ref.tobytes()
