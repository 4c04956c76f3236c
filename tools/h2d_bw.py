#!/usr/bin/env python
"""Host<->device copy bandwidth of the box from page-locked memory, to put the end-to-end numbers in context.
  python tools/h2d_bw.py                 one GPU, H2D and D2H
  python tools/h2d_bw.py --gpus 8        one process per GPU (torchrun), all copying AT THE SAME TIME, with and without
                                         binding each process to its GPU's NUMA node (bench.py: numa_bind / h2d_ceiling)
Prints one JSON line (rank 0)."""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--mb", type=int, default=256)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 and a.gpus > 1:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29531", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    import torch
    import bench

    class A:
        pass

    out = {}
    # first unpinned (the process may run anywhere), then bound to the GPU's node: page-locked buffers are allocated inside
    # h2d_ceiling, i.e. after the binding
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    saved, bench.numa_bind = bench.numa_bind, (lambda t, l: {"bound": False, "why": "unpinned leg"})
    ctx = bench.Ctx(A())
    out["unpinned"] = bench.h2d_ceiling(ctx, mb=a.mb)
    bench.numa_bind = saved
    ctx.numa = bench.numa_bind(torch, local)
    out["numa_bound"] = bench.h2d_ceiling(ctx, mb=a.mb)
    out["numa"] = ctx.numa
    # D2H, one direction at a time
    n = a.mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.ones(n, dtype=torch.uint8, device=ctx.dev)
    h.copy_(d, non_blocking=True)
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(6):
        h.copy_(d, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    out["d2h_gbs_per_gpu"] = 6 * n / (ctx.reduce(e0.elapsed_time(e1), "MAX") * 1e-3) / 1e9
    if ctx.rank == 0:
        print(json.dumps(out))
    if ctx.dist is not None:
        ctx.dist.destroy_process_group()


if __name__ == "__main__":
    main()
