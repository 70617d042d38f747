"""torchrun --nproc-per-node R scripts/strips_check_nccl.py : R real ranks (NCCL) against the undivided map on rank 0."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch, torch.distributed as dist
import soillib_b200 as soil
import oracle_api as orc
from soillib_b200.strips import DistTransport, LocalTransport, Strip, erode_strips, partition_rows

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nx, ny, P, steps = 512, 384, 4096, 4
h = torch.from_numpy(orc.noise(nx, ny, 3.0))
p = soil.param_t(); C.memmove(C.byref(p._c), C.byref(orc.example_param()), 112)
scale = (20.0 / nx, 20.0 / ny, 4.0)
lo, hi = partition_rows(nx, world)[rank]
halo = int(sys.argv[sys.argv.index("--halo") + 1]) if "--halo" in sys.argv else 1
s = Strip(nx, ny, lo, hi, scale, P, albedo=True, halo=halo, view="--view" in sys.argv)
s.set_from_global(bedrock=h, uplift=h.clamp(0, 1))
tr = DistTransport()
if "--peers" in sys.argv:
    tr.connect([s])
overlap = "--no-overlap" not in sys.argv
rounds = erode_strips([s], tr, p, steps, mass_sum=True, overlap=overlap)
mine = s.own(s.layers[s.cur]).contiguous()
sums = s.mass_sum.clone()
dist.all_reduce(sums)
parts = [torch.empty((b - a, ny, 2), device="cuda") for a, b in partition_rows(nx, world)]
dist.all_gather(parts, mine) if len(set(t.shape for t in parts)) == 1 else None
if len(set(t.shape for t in parts)) != 1:
    for r, t in enumerate(parts):
        if r == rank: t.copy_(mine)
        dist.broadcast(t, r)
if rank == 0:
    got = torch.cat(parts).cpu().numpy()
    w = Strip(nx, ny, 0, nx, scale, P, albedo=True)
    w.set_from_global(bedrock=h, uplift=h.clamp(0, 1))
    erode_strips([w], LocalTransport(), p, steps)
    want = w.layers[w.cur].cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(sums.cpu().numpy(), got.astype(np.float64).sum(axis=(0, 1)), rtol=1e-12)
    print(f"NCCL strips ok (halo {halo}, {'peer-mapped' if '--peers' in sys.argv else 'hand-off'}, overlap {overlap}, view {'--view' in sys.argv}): {world} ranks == whole map "
          f"within 2e-6 after {steps} steps; hand-off rounds last step {rounds}; max|diff| {np.abs(got - want).max():.3e}; mass sums ok", flush=True)
dist.barrier()
dist.destroy_process_group()
