"""phase clocks of k_tile from a -DSOILB200_TILE_TIMING build (SOILB200_LIB=build/lib_timing.so)"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import soillib_b200 as soil
from soillib_b200 import silt
n = 8192
p = soil.noise_t(); p.seed, p.ext = 3.0, (float(n), float(n))
h = soil.noise(silt.shape(n, n), p, host=silt.gpu)
rain = silt.tensor.from_torch(torch.ones((n, n), device="cuda"))
lib = soil._load_library()
buf = (C.c_ulonglong * 16)()
for name, g in (("steepest", soil.steepest(h, soil.d8)), ("gibbs", soil.random_weighted(h, soil.d8, 0, 0, 10.0))):
    soil.accumulate(g, rain, soil.d8); torch.cuda.synchronize()
    lib.soilb200_debug_tile_clocks(buf)
    soil.accumulate(g, rain, soil.d8); torch.cuda.synchronize()
    lib.soilb200_debug_tile_clocks(buf)
    tiles = (n // 64) ** 2
    v = [x / tiles for x in buf]
    print(name, "cycles per tile: load_graph %.0f  src %.0f  perimeter %.0f  queue %.0f  rounds %.0f  write %.0f | dense rounds %.1f (%.0f cycles) sparse rounds %.1f (%.0f cycles)" %
          (v[1], 0, v[2], v[3], v[4], v[5], v[8], v[10], v[9], v[11]))
