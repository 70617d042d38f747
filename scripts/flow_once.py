"""each stencil op of the flow path twice at a given size - the ncu target for the per-kernel captures"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import soillib_b200 as soil
from soillib_b200 import silt

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
p = soil.noise_t(); p.seed, p.ext = 3.0, (float(n), float(n))
h = soil.noise(silt.shape(n, n), p, host=silt.gpu)
scale = (20.0 / n, 20.0 / n, 4.0)
rain = silt.tensor.from_torch(torch.ones((n, n), device="cuda"))
L = torch.rand((n, n, 2), device="cuda") * 0.05
D = torch.zeros_like(L)
prm = soil.param_t()
for rep in range(2):
    soil.direction(h, soil.d8)
    soil.steepest(h, soil.d8)
    soil.normal(h, scale)
    soil.mass_creep(silt.tensor.from_torch(D), silt.tensor.from_torch(L), scale, prm)
    g = soil.random_weighted(h, soil.d8, 0, rep, 10.0)
    soil.accumulate(g, rain, soil.d8)
torch.cuda.synchronize()
print("ok")
