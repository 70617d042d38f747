"""torchrun --nproc-per-node R scripts/accumulate_strips_nccl.py [rows_per_rank] [ny]: steepest + accumulate over R row
strips on R GPUs (NCCL all-gather of the boundary summaries) against the undivided map on rank 0; device times."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import soillib_b200 as soil
from soillib_b200 import silt
from soillib_b200.accumulate_strips import accumulate_strip_dist
from soillib_b200.flow_strips import HeightStrip
from soillib_b200.strips import DistTransport, partition_rows

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ny = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
nx = rows * world
p = soil.noise_t(); p.seed, p.ext = 3.0, (float(ny), float(ny))
full = soil.noise(silt.shape(nx, ny), p, host=silt.gpu).data            # every rank synthesises the map, keeps its strip
lo, hi = partition_rows(nx, world)[rank]
strip = HeightStrip(nx, ny, lo, hi, halo=2)
strip.own(strip.block).copy_(full[lo:hi])
strip.refresh(DistTransport())                                           # halo rows from the neighbours over NCCL
rain = torch.ones((hi - lo, ny), device="cuda")

def step():
    g = strip.steepest(soil.d8).data
    return accumulate_strip_dist(g, rain, lo, hi, nx, soil.d8)

acc = step(); torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); acc = step(); e1.record(); torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1)], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
parts = [torch.empty((b - a, ny), device="cuda") for a, b in partition_rows(nx, world)]
dist.all_gather(parts, acc)
if rank == 0:
    hd = silt.tensor.from_torch(full)
    W = silt.tensor.from_torch
    def whole():
        return soil.accumulate(soil.steepest(hd, soil.d8), W(torch.ones((nx, ny), device="cuda")), soil.d8).data
    want = whole(); torch.cuda.synchronize()
    e0.record(); want = whole(); e1.record(); torch.cuda.synchronize()
    ok = torch.equal(torch.cat(parts), want)
    print(f"accumulate over {world} strips of {rows}x{ny}: {'EQUAL to' if ok else 'DIFFERS from'} the undivided map; "
          f"{t.item():.2f} ms on {world} GPUs, {e0.elapsed_time(e1):.2f} ms on one; max accumulation {want.max().item():.0f}", flush=True)
dist.barrier(); dist.destroy_process_group()
