"""flow ops at a given size: times ours and (optionally) the reference kernels; used under ncu too."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import soillib_b200 as soil
from soillib_b200 import silt

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
with_ref = "--ref" in sys.argv
reps = 20
p = soil.noise_t(); p.seed, p.ext = 3.0, (float(n), float(n))
h = soil.noise(silt.shape(n, n), p, host=silt.gpu)
rain = silt.tensor.from_torch(torch.ones((n, n), device="cuda"))
scale = (20.0 / n, 20.0 / n, 4.0)

def timeit(name, fn, nbytes):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{name:28s} {ms:9.3f} ms  {nbytes * n * n / ms / 1e6:8.1f} GB/s algorithmic", flush=True)

timeit("direction d8", lambda: soil.direction(h, soil.d8), 8)
timeit("steepest d8", lambda: soil.steepest(h, soil.d8), 8)
timeit("random_weighted d8 (resume)", lambda: soil.random_weighted(h, soil.d8, 0, timeit.k.__next__(), 10.0) if False else None, 8) if False else None
k = [0]
def rw():
    soil.random_weighted(h, soil.d8, 0, k[0], 10.0); k[0] += 1
timeit("random_weighted d8", rw, 8)
g = soil.steepest(h, soil.d8)
timeit("accumulate d8 (steepest)", lambda: soil.accumulate(g, rain, soil.d8), 12)
grw = soil.random_weighted(h, soil.d8, 0, 0, 10.0)
timeit("accumulate d8 (random)", lambda: soil.accumulate(grw, rain, soil.d8), 12)
timeit("normal", lambda: soil.normal(h, scale), 16)
L = torch.rand((n, n, 2), device="cuda") * 0.05
D = torch.zeros_like(L)
prm = soil.param_t()
timeit("mass_creep", lambda: soil.mass_creep(silt.tensor.from_torch(D), silt.tensor.from_torch(L), scale, prm), 24)
s2 = (20.0 / n, 20.0 / n)
timeit("gradient", lambda: soil.gradient(h, s2), 12)
timeit("negslope", lambda: soil.negslope(h, s2), 8)
h1 = silt.tensor.from_torch(h.data.unsqueeze(-1).contiguous())
timeit("laplacian (1 ch)", lambda: soil.laplacian(h1, s2), 8)
timeit("slope", lambda: soil.slope(h, g, s2), 12)
hb = silt.tensor.from_torch(h.data.unsqueeze(-1).clone())
timeit("gaussian_blur (1 ch)", lambda: soil.gaussian_blur(hb, 2.5), 16)
if with_ref:
    import ref_api as ref
    timeit("REF direction d8", lambda: ref.direction(h.data, 1), 8)
    timeit("REF random_weighted d8", lambda: ref.random_weighted(h.data, 1, 0, 3, 10.0), 8)
    timeit("REF accumulate d8 (steepest)", lambda: ref.accumulate(g.data, rain.data, None, 1), 12)
    import ctypes as C, oracle_api as orc
    op = orc.default_param()
    timeit("REF mass_creep", lambda: ref.mass_creep(D, L, scale, op), 24)
    timeit("REF gradient", lambda: ref.gradient(h.data, s2), 12)
    timeit("REF negslope", lambda: ref.negslope(h.data, s2), 8)
    timeit("REF laplacian (1 ch)", lambda: ref.laplacian(h1.data, s2), 8)
    timeit("REF slope", lambda: ref.slope(h.data, g.data, s2), 12)
    hb2 = h.data.unsqueeze(-1).clone()
    timeit("REF gaussian_blur (1 ch)", lambda: ref.lib().ref_gaussian_blur(ref._p(hb2), n, n, 1, C.c_float(2.5)), 16)
