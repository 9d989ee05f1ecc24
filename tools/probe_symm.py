"""Probe (multi-GPU box): does torch symmetric memory work here, is multicast (NVLS) available, and how do its
all-reduce variants compare with NCCL on the J^T J v payload (59 P floats)?   torchrun --nproc-per-node N tools/probe_symm.py"""
import datetime
import os
import sys
import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=90))
rank, world = dist.get_rank(), dist.get_world_size()
n = 59_000_000


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


x = torch.randn(n, device=dev)
ms = timed(lambda: dist.all_reduce(x))
if rank == 0:
    print(f"world {world}: NCCL all_reduce 236 MB: {ms:.3f} ms", flush=True)
try:
    import torch.distributed._symmetric_memory as sm
    t = sm.empty(n, dtype=torch.float32, device=dev)
    hdl = sm.rendezvous(t, dist.group.WORLD.group_name)
    if rank == 0:
        print("symm_mem rendezvous ok; multicast_ptr =", getattr(hdl, "multicast_ptr", None), "buffer_ptrs", len(hdl.buffer_ptrs),
              "signal_pad_ptrs", len(hdl.signal_pad_ptrs), flush=True)
    t.copy_(torch.randn(n, device=dev))
    ops = torch.ops.symm_mem
    for name in ("multimem_all_reduce_", "two_shot_all_reduce_", "one_shot_all_reduce"):
        try:
            f = getattr(ops, name)
            ms = timed(lambda: f(t, "sum", dist.group.WORLD.group_name))
            if rank == 0:
                print(f"  symm_mem.{name}: {ms:.3f} ms", flush=True)
        except Exception as e:     # noqa: BLE001
            if rank == 0:
                print(f"  symm_mem.{name}: FAILED {type(e).__name__}: {str(e)[:200]}", flush=True)
except Exception as e:             # noqa: BLE001
    if rank == 0:
        print("symmetric memory unavailable:", type(e).__name__, str(e)[:300], flush=True)
dist.barrier()
dist.destroy_process_group()
