"""torchrun worker: point-sharded multi-GPU LM solve must reproduce the single-GPU solve.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/multi_gpu_worker.py [shape]
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba  # noqa: E402


def main():
    shapes = (sys.argv[1] if len(sys.argv) > 1 else "ladybug49").split("+")
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    # "a+b": several problems in ONE process (b with more cameras: the peer-memory exchange block is rebuilt)
    flags = [solve_one(shape, dev, rank, world) for shape in shapes]
    flag = torch.tensor([int(all(flags))], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if rank == 0:
        print("MULTI_GPU_OK" if flag.item() == 1 else "MULTI_GPU_FAIL", flush=True)
    sys.exit(0 if flag.item() == 1 else 1)


def solve_one(shape, dev, rank, world):
    gap = shape.endswith("_gap")
    sc = b200ba.synth.make_ba_scene(shape[:-4] if gap else shape, seed=0)
    if gap:
        # camera 0 sees only points of the first half: the rank that owns the second half has NO row of it and
        # still has to publish a zero partial sum for it every iteration
        deg = np.bincount(sc.pt_idx, minlength=sc.P)
        drop = (sc.cam_idx == 0) & (sc.pt_idx >= sc.P // 2) & (deg[sc.pt_idx] > 2)
        sc.cam_idx, sc.pt_idx, sc.obs = sc.cam_idx[~drop], sc.pt_idx[~drop], sc.obs[~drop]
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, device=dev, distributed=True)
    assert ba.world == world
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    params, st = ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None)
    ok = True
    if rank == 0:
        single = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, device=dev, distributed=False)
        p1, s1 = single.optimize(x0, single.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None)
        dl = np.abs(st.loss_trajectory[:min(len(s1.loss_trajectory), len(st.loss_trajectory))] -
                    s1.loss_trajectory[:min(len(s1.loss_trajectory), len(st.loss_trajectory))]) / s1.loss_trajectory[:min(len(s1.loss_trajectory), len(st.loss_trajectory))]
        print(f"world {world} iters {st.iter_num} vs {s1.iter_num}; max rel loss dev {dl.max():.3e}; "
              f"max |dparam| {np.abs(params - p1).max():.3e}", flush=True)
        ok = st.iter_num == s1.iter_num and dl.max() < 1e-5 and not np.isnan(params).any()
    # every rank must hold the same full parameter vector
    t = torch.from_numpy(params).to(dev)
    ref = t.clone()
    dist.broadcast(ref, 0)
    same = bool(torch.equal(t, ref))
    return ok and same


if __name__ == "__main__":
    main()
