"""accumulate / accumulate_decay at a given size: ours (tiled and global walk) and the reference's rake-compress,
checked against each other.  usage: acc_probe.py [size] [--ref]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import soillib_b200 as soil
from soillib_b200 import silt

n = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 8192
p = soil.noise_t(); p.seed, p.ext = 3.0, (float(n), float(n))
h = soil.noise(silt.shape(n, n), p, host=silt.gpu)
rain = silt.tensor.from_torch(torch.ones((n, n), device="cuda"))
src = silt.tensor.from_torch(h.data.abs() + 0.25)
dec = silt.tensor.from_torch(torch.full((n, n), 0.97, device="cuda"))


def timeit(name, fn, nbytes, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{name:40s} {ms:9.3f} ms  {nbytes * n * n / ms / 1e6:8.1f} GB/s algorithmic", flush=True)


g = soil.steepest(h, soil.d8)
grw = soil.random_weighted(h, soil.d8, 0, 0, 10.0)
timeit("accumulate d8 (steepest graph)", lambda: soil.accumulate(g, rain, soil.d8), 12)
timeit("accumulate d8 (Gibbs-random graph)", lambda: soil.accumulate(grw, rain, soil.d8), 12)
timeit("accumulate_decay d8 (steepest graph)", lambda: soil.accumulate_decay(g, src, dec, soil.d8), 16)
timeit("accumulate_decay d8 (Gibbs-random graph)", lambda: soil.accumulate_decay(grw, src, dec, soil.d8), 16)
if "--ref" in sys.argv:
    import ref_api as ref
    timeit("REF accumulate d8 (steepest)", lambda: ref.accumulate(g.data, rain.data, None, 1), 12, reps=3)
    timeit("REF accumulate_decay d8 (steepest)", lambda: ref.accumulate(g.data, src.data, dec.data, 1), 16, reps=3)
    a, b = soil.accumulate(g, rain, soil.d8).data, ref.accumulate(g.data, rain.data, None, 1)
    print("unit rain vs reference: identical" if torch.equal(a, b) else f"unit rain vs reference: {(a != b).sum().item()} cells differ")
    a, b = soil.accumulate_decay(g, src, dec, soil.d8).data, ref.accumulate(g.data, src.data, dec.data, 1)
    rel = ((a - b).abs() / b.abs().clamp_min(1e-30)).max().item()
    print(f"decay vs reference: max relative difference {rel:.3e}")
