#!/usr/bin/env python
"""Times the one collective of the path: allreduce(sum, fp32) of the packed 59 P-float gradient buffer (236 MB at
P = 1M) over the ranks of one box, as LinearSolverFunctions.normal_matvec issues it once per CG iteration.
    torchrun --nproc-per-node N tools/bench_allreduce.py [--points 1000000] [--iters 20]
Prints one JSON line on rank 0 (device-timed, max over ranks)."""
import argparse
import json
import os

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=1000000)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    n = 59 * a.points
    x = torch.ones(n, dtype=torch.float32, device=dev)
    for _ in range(5):
        dist.all_reduce(x)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        dist.all_reduce(x)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if dist.get_rank() == 0:
        w = dist.get_world_size()
        print(json.dumps({"world": w, "bytes": 4 * n, "ms": float(ms), "algbw_GBps": 4 * n / float(ms) / 1e6,
                          "busbw_GBps": 4 * n / float(ms) / 1e6 * 2 * (w - 1) / w,
                          "env": {k: v for k, v in os.environ.items() if k.startswith("NCCL_")}}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
