"""direction / steepest at 8192^2 with the variant SOILB200_STEEPEST selects (default: per-warp cp.async ring; staged:
1-D bulk copies; tma: 2-D tensor-map boxes).  One process per variant: the variable is read once."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import soillib_b200 as soil
from soillib_b200 import silt

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
p = soil.noise_t(); p.seed, p.ext = 3.0, (float(n), float(n))
h = soil.noise(silt.shape(n, n), p, host=silt.gpu)
for name, fn in (("direction d8", lambda: soil.direction(h, soil.d8)), ("steepest d8", lambda: soil.steepest(h, soil.d8)),
                 ("direction d4", lambda: soil.direction(h, soil.d4))):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 30
    print(f"{os.environ.get('SOILB200_STEEPEST', 'default'):8s} {name:14s} {n}^2 {ms:8.4f} ms  {8 * n * n / ms / 1e6:8.1f} GB/s algorithmic  "
          f"{8 * n * n / ms / 1e6 / 6542.4:.3f} of the measured HBM peak", flush=True)
