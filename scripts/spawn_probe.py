"""one GPU playing rank 0 of `world` ranks (no peers): what do k_spawn and the N-times-larger transport grid cost?
usage: spawn_probe.py [world ...]   (run under ncu --metrics gpu__time_duration.sum for the per-kernel times)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import soillib_b200 as soil
from soillib_b200 import silt
from soillib_b200.strips import LocalTransport, Strip, erode_strips, partition_rows

n = 8192
for world in [int(a) for a in sys.argv[1:]] or [1, 2, 8]:
    nx, P = n * world, 32768 * world
    npar = soil.noise_t(); npar.seed, npar.ext = 3.0, (float(n), float(n))
    blk = soil.noise(silt.shape(n + 256, n), npar, host=silt.gpu).data      # rank 0's rows + its halo
    lo, hi = partition_rows(nx, world)[0]
    s = Strip(nx, n, lo, hi, (20.0 / n, 20.0 / n, 4.0), P, albedo=True, halo=256, view=True)
    s.set_from_global(bedrock=blk, uplift=blk.clamp(0.0, 1.0))     # the strip slices its rows [0, rows) out of it
    p = soil.param_t()
    p.timeStep, p.maxage, p.uplift, p.evapRate, p.fluvialExponent, p.exitSlope = 10.0, 512, 0.01, 0.0005, 0.01, 0.025
    tr = LocalTransport()
    tr.connect([s])
    erode_strips([s], tr, p, 3, export=False)
    torch.cuda.synchronize()
    marks = []
    erode_strips([s], tr, p, 10, events=marks, export=False)
    torch.cuda.synchronize()
    seg = [sum(marks[6 * i + j].elapsed_time(marks[6 * i + j + 1]) for i in range(10)) / 10 for j in range(5)]
    print(f"world {world}: particles {seg[0]:.3f} ms  dense {seg[2]:.3f} ms", flush=True)
    del s, blk
    torch.cuda.empty_cache()
