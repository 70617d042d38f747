"""Particle-phase time of the erosion step against the particle count (8192^2, config C3's parameters).

Separates the two possible bounds of k_transport_both: time proportional to P = a throughput limit (DRAM random
access / LSU wavefronts), flat = the latency of one particle's dependent chain.  usage: particle_sweep.py [size] [P ...]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import soillib_b200 as soil

size = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
Ps = [int(v) for v in sys.argv[2:]] or [8192, 16384, 32768, 65536, 131072]
lib = soil._load_library()
for P in Ps:
    model, data, track, p, h = bench.build_workload(soil, size, size, P, True)
    for _ in range(4):
        soil.erode(model, data, track, p, 1)
    torch.cuda.synchronize()
    lib.soilb200_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    soil.erode(model, data, track, p, 10)
    e1.record()
    torch.cuda.synchronize()
    pm, dm, st, hot = C.c_double(), C.c_double(), C.c_int64(), C.c_uint64()
    lib.soilb200_profile_read(C.byref(pm), C.byref(dm), C.byref(st), C.byref(hot))
    lib.soilb200_profile_enable(0)
    print(f"P={P:7d} size={size} step {e0.elapsed_time(e1) / 10:.3f} ms  particles {pm.value / st.value:.3f} ms  dense {dm.value / st.value:.3f} ms  "
          f"cells touched/step {hot.value / st.value:.0f}", flush=True)
    del model, data, track
    torch.cuda.empty_cache()
