"""Replica-parallel multi-flow accumulation under torchrun (one process per GPU, NCCL):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/multiflow_replicas.py [size] [K]
Every rank builds the same DEM, takes its share of the K samples, one all-reduce forms the mean.  Rank 0 also
computes all K samples alone and reports the difference and both device times (max over ranks)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import soillib_b200 as soil
from soillib_b200 import silt

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
K = int(sys.argv[2]) if len(sys.argv) > 2 else 16
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("nccl")
p = soil.noise_t(); p.seed, p.ext = 3.0, (float(n), float(n))
h = soil.noise(silt.shape(n, n), p, host=silt.gpu)
rain = silt.tensor.from_torch(torch.ones((n, n), device="cuda"))

def timed(fn, collective=True):
    """device time of fn after one warm-up call; with collective=True every rank must call it (max over ranks)"""
    fn(); torch.cuda.synchronize()
    if world > 1 and collective: dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); out = fn(); e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1 and collective: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return out, t.item()

if world > 1:
    mean, ms = timed(lambda: soil.multiflow_replicas(h, rain, soil.d8, K, 10.0))
else:
    mean, ms = timed(lambda: soil.multiflow(h, rain, soil.d8, K, 10.0))
if rank == 0:
    solo, ms1 = timed(lambda: soil.multiflow(h, rain, soil.d8, K, 10.0), collective=False) if world > 1 else (mean, ms)
    err = ((mean.data - solo.data).abs() / solo.data.abs().clamp_min(1e-30)).max().item()
    print(f"multiflow {n}x{n} K={K}: {world} GPU(s) {ms:.2f} ms ({K * n * n / ms / 1e6:.1f} Gcell-samples/s), "
          f"1 GPU {ms1:.2f} ms, speed-up {ms1 / ms:.2f}, max rel diff vs one GPU {err:.2e}", flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
