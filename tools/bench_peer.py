"""Multi-GPU microbenchmark of the parameter-space reduction (59 P floats): this library's multimem kernel over NVSwitch
peer memory (csrc/peer_reduce.cu, with and without its two barriers) against NCCL's allreduce.
    torchrun --nproc-per-node N tools/bench_peer.py [P]"""
import datetime
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting-lm_b200"))
import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=90))
rank, world = dist.get_rank(), dist.get_world_size()
P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
n = 59 * P
from diff_gaussian_rasterization import _C
from solver.solver_functions import _PeerAccumulator


def timed(fn, iters=20):
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
t_nccl = timed(lambda: dist.all_reduce(x))
pa = _PeerAccumulator(n, dev, dist)
# correctness: sum of rank-dependent vectors
g = torch.Generator(device=dev).manual_seed(7)
base = torch.randn(n, device=dev, generator=g)
acc = pa.view(n)
n_pad = (n + 3) // 4 * 4
acc.copy_(base * (rank + 1))
pa.allreduce(acc)
want = base * (world * (world + 1) / 2)
err = float((acc - want).abs().max() / want.abs().max())
t_full = timed(lambda: pa.allreduce(acc))
t_kernel = timed(lambda: _C.multimem_allreduce(pa.mc, n_pad, rank, world))
t_barrier = timed(lambda: pa.hdl.barrier(channel=0))
if rank == 0:
    mb = 4 * n / 1e6
    print(f"world {world} payload {mb:.0f} MB unroll {os.environ.get('GSLM_MM_UNROLL', '4')} ctas/SM {os.environ.get('GSLM_MM_CTAS', '4')}: "
          f"NCCL allreduce {t_nccl:.3f} ms | multimem barrier+kernel+barrier {t_full:.3f} ms "
          f"(kernel alone {t_kernel:.3f} ms, one barrier {t_barrier:.3f} ms) | max rel err {err:.2e}", flush=True)
dist.barrier()
dist.destroy_process_group()
