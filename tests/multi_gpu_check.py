"""Multi-GPU check of the C ABI's group entry points (not a pytest module: the GPU test tier has one GPU).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \\
        --master-port 29611 tests/multi_gpu_check.py

One process per GPU; torch.distributed only carries the group id.  Every rank calls ag2_group_detect /
ag2_group_detect_importance (csrc/group.cu: peer-memory windows over NVLink, NCCL as the fallback); rank 0
additionally runs the same jobs on its GPU alone and compares byte for byte:
  (1) detect with host inputs, (2) detect with device-resident inputs, repeated (steady state),
  (3) importance-sampling rounds without a top-k bound (unbounded lists),
  (4) importance-sampling rounds with LeNet scores, a top-k per round and the final clustering,
  (5) (1) again on the NCCL transport (AG2_GROUP_NCCL=1)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from agile_grasp2_b200 import capi, scenes  # noqa: E402
from common import make_params  # noqa: E402


def make_group(p, rank, world, local, dev):
    uid = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        uid.copy_(torch.frombuffer(bytearray(capi.group_unique_id()), dtype=torch.uint8))
    dist.broadcast(uid, src=0)
    return capi.Group(p, uid=uid.cpu().numpy().tobytes(), rank=rank, world=world, device=local)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    sc = scenes.make_scene("tiny", num_samples=90)
    n, s = len(sc["xyz"]), len(sc["samples"])
    ok = True

    def report(what, same, extra=""):
        nonlocal ok
        if rank == 0:
            print("%-64s %s %s" % (what, "equal to one GPU" if same else "DIFFERS from one GPU", extra), flush=True)
            ok &= bool(same)

    # (1), (2) plain detect
    p = make_params(capi.MODE_PREDICTION, min_score_diff=-1e9, num_selected=25)
    want = None
    if rank == 0:
        with capi.Context(p, device=local) as one:
            one.init_lenet_xavier(3)
            one.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); one.set_sample_indices(sc["samples"])
            one.detect()
            want = one.hypotheses(capi.LIST_SELECTED).tobytes()
    for transport in ("peer", "nccl"):
        os.environ["AG2_GROUP_NCCL"] = "1" if transport == "nccl" else "0"
        grp = make_group(p, rank, world, local, dev)
        grp.init_lenet_xavier(3)
        if rank == 0:
            got = grp.detect(sc["xyz"], sc["cam_mask"], sc["num_cams"], sc["samples"])
        else:
            got = grp.detect(None, None, sc["num_cams"], None, n=n, n_samples=s)
        report("detect, host inputs, %d GPUs, transport asked %s / peer stores %s" % (world, transport, grp.peer_stores()),
               rank != 0 or (got.tobytes() == want and len(got) == 25))
        if rank == 0:
            xd, cd = torch.from_numpy(sc["xyz"]).to(dev), torch.from_numpy(sc["cam_mask"]).to(dev)
            sd = torch.from_numpy(np.asarray(sc["samples"], np.int32)).to(dev)
            torch.cuda.synchronize()
            a3 = (xd.data_ptr(), cd.data_ptr(), sd.data_ptr())
        else:
            a3 = (0, 0, 0)
        same = True
        for _ in range(3):
            got = grp.detect_ptr(a3[0], n, a3[1], sc["num_cams"], a3[2], s, True)
            same &= rank != 0 or got.tobytes() == want
        report("detect, device-resident inputs x3 (%s)" % transport, same)
        grp.close()
    os.environ["AG2_GROUP_NCCL"] = "0"

    # (3) importance-sampling rounds, unbounded lists
    p2 = make_params(capi.MODE_NONE, num_selected=10)
    grp = make_group(p2, rank, world, local, dev)
    ip = grp.local_ctx(0).default_is_params()
    ip.num_iterations, ip.num_samples, ip.seed = 3, 50, 21
    args = (sc["xyz"], sc["cam_mask"], sc["num_cams"], sc["samples"]) if rank == 0 else (None, None, sc["num_cams"], None)
    got2 = grp.detect_importance(*args, ip, n=n, n_samples=s)
    if rank == 0:
        with capi.Context(p2, device=local) as one:
            one.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); one.set_sample_indices(sc["samples"])
            one.detect_importance(ip)
            want2 = one.hypotheses(capi.LIST_IMPORTANCE)
        report("importance sampling, mode 0, %d rounds" % (ip.num_iterations + 1),
               got2.tobytes() == want2.tobytes() and len(got2) > 10, "(%d hands)" % len(got2))
    grp.close()

    # (4) importance sampling with scores, top-k per round, clusters at the end
    p3 = make_params(capi.MODE_PREDICTION, min_score_diff=-1e9, num_selected=30)
    p3.min_inliers = 2
    grp = make_group(p3, rank, world, local, dev)
    grp.init_lenet_xavier(3)
    got3 = grp.detect_importance(*args, ip, n=n, n_samples=s)
    if rank == 0:
        with capi.Context(p3, device=local) as one:
            one.init_lenet_xavier(3)
            one.set_cloud(sc["xyz"], sc["cam_mask"], sc["num_cams"]); one.set_sample_indices(sc["samples"])
            one.detect_importance(ip)
            want3 = one.hypotheses(capi.LIST_CLUSTERS)
        report("importance sampling, mode 1 + clusters", got3.tobytes() == want3.tobytes() and len(got3) > 0,
               "(%d clusters)" % len(got3))
    grp.close()

    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, src=0)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MULTI_GPU_CHECK", "OK" if ok else "FAILED")
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
