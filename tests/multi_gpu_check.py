"""Multi-GPU check over NCCL (not a pytest module: the GPU test tier has one GPU).  Launch:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29611 tests/multi_gpu_check.py

Every rank drives its own GPU through libag2; rank 0 additionally runs the same jobs on its GPU alone
and compares: (1) the sharded detect call (packed broadcast, sharded normals, device all-gather) and
(2) the sharded importance-sampling rounds must give the single-GPU lists byte for byte."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from agile_grasp2_b200 import capi, scenes, sharded  # noqa: E402
from common import make_params  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    sc = scenes.make_scene("tiny", num_samples=90)
    n, s = len(sc["xyz"]), len(sc["samples"])
    ok = True

    # (1) plain detect, GPU fast path
    p = make_params(capi.MODE_PREDICTION, min_score_diff=-1e9, num_selected=25)
    ctx = capi.Context(p, device=local)
    ctx.init_lenet_xavier(3)
    det = sharded.ShardedDetector(ctx, rank=rank, world=world, device=dev)
    if rank == 0:
        packed = sharded.pack_inputs(sc["xyz"], sc["cam_mask"], sc["samples"], dev)[0]
        torch.cuda.synchronize()
    else:
        packed = torch.empty(12 * n + 4 * s + n, dtype=torch.uint8, device=dev)
    got = det.detect_packed(packed, n, s, sc["num_cams"])
    if rank == 0:
        with capi.Context(p, device=local) as one:
            one.init_lenet_xavier(3)
            one.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); one.set_sample_indices(sc["samples"])
            one.detect()
            want = one.hypotheses(capi.LIST_SELECTED)
        same = got.tobytes() == want.tobytes()
        print("detect_packed over %d GPUs: %d selected, equal to one GPU: %s" % (world, len(got), same))
        ok &= same and len(got) == 25

    # (2) importance-sampling rounds
    p2 = make_params(capi.MODE_NONE, num_selected=10)
    ctx2 = capi.Context(p2, device=local)
    ip = ctx2.default_is_params()
    ip.num_iterations, ip.num_samples, ip.seed = 3, 50, 21
    det2 = sharded.ShardedDetector(ctx2, rank=rank, world=world, device=dev)
    args = (sc["xyz"], sc["cam_mask"], sc["num_cams"], sc["samples"]) if rank == 0 else (None, None, 0, None)
    got2 = det2.detect_importance_numpy(*args, ip)
    if rank == 0:
        with capi.Context(p2, device=local) as one:
            one.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); one.set_sample_indices(sc["samples"])
            one.detect_importance(ip)
            want2 = one.hypotheses(capi.LIST_IMPORTANCE)
        same = got2.tobytes() == want2.tobytes()
        print("importance sampling over %d GPUs: %d hands in %d rounds, equal to one GPU: %s"
              % (world, len(got2), ip.num_iterations + 1, same))
        ok &= same and len(got2) > 10
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, src=0)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MULTI_GPU_CHECK", "OK" if ok else "FAILED")
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
