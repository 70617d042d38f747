"""one GPU: what does the strip machinery itself cost the particle phase? (whole-map strip with / without a peer table,
two logical strips on one device)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import soillib_b200 as soil
from soillib_b200 import silt
from soillib_b200.strips import LocalTransport, Strip, erode_strips, partition_rows

n = 8192
def run(world, peers, label, steps=10):
    nx = n * world
    P = 32768 * world
    npar = soil.noise_t(); npar.seed, npar.ext = 3.0, (float(n), float(n))
    full = soil.noise(silt.shape(nx, n), npar, host=silt.gpu).data
    strips = []
    for lo, hi in partition_rows(nx, world):
        s = Strip(nx, n, lo, hi, (20.0 / n, 20.0 / n, 4.0), P, albedo=True, halo=256, view=True)
        s.set_from_global(bedrock=full, uplift=full.clamp(0, 1))
        strips.append(s)
    del full
    p = soil.param_t()
    p.timeStep, p.maxage, p.uplift, p.evapRate, p.fluvialExponent, p.exitSlope = 10.0, 512, 0.01, 0.0005, 0.01, 0.025
    tr = LocalTransport()
    if peers: tr.connect(strips)
    erode_strips(strips, tr, p, 3, export=False)
    torch.cuda.synchronize()
    marks = [] if peers else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); erode_strips(strips, tr, p, steps, events=marks, export=False); e1.record(); torch.cuda.synchronize()
    msg = f"{label}: {e0.elapsed_time(e1) / steps:.3f} ms/step"
    if marks:
        seg = [sum(marks[6 * i + j].elapsed_time(marks[6 * i + j + 1]) for i in range(steps)) / steps for j in range(5)]
        msg += "  particles %.3f  dense %.3f" % (seg[0], seg[2])
    print(msg, flush=True)

run(1, False, "whole map, no peer table")
run(1, True, "whole map, peer table of one")
run(2, True, "two strips of 8192 rows on one GPU (both phases run back to back)")
