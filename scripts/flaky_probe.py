"""which side of test_erosion_vs_reference_kernels varies on the first run of a fresh process?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import oracle_api as oracle, ref_api as ref
import soillib_b200 as soil
from test_gpu_erosion import _run_fused
nx, ny, P, steps = 128, 96, 1024, 3
op = oracle.default_param(maxage=128) if "default" in sys.argv else oracle.example_param()
h = oracle.noise(nx, ny, 3.0)
scale = (20.0 / nx, 20.0 / ny, 4.0)
def make():
    m = oracle.HostModel(h, scale, P, albedo=True, seed=0, uplift=np.clip(h, 0, 1)); m.layers[..., 1] = 0.01; return m
order = sys.argv[1] if len(sys.argv) > 1 else "ours"
res = {}
def ours(tag):
    o, _, _ = _run_fused(soil, oracle, make(), op, steps, True, True); res[tag] = o["waterHeight"].cpu().numpy()
def theirs(tag):
    dm = ref.DeviceModel(make())
    for _ in range(steps): dm.step(op, True, nofma=True)
    torch.cuda.synchronize(); res[tag] = dm.get("waterHeight")
seq = [("o1", ours), ("r1", theirs), ("o2", ours), ("r2", theirs)] if order == "ours" else [("r1", theirs), ("o1", ours), ("r2", theirs), ("o2", ours)]
for t, f in seq: f(t)
def bad(a, b):
    fin = np.isfinite(a) & np.isfinite(b)
    return int((np.abs(a[fin] - b[fin]) > 1e-4 * np.abs(b[fin]) + 1e-9).sum())
print(order, "o1-o2", bad(res["o1"], res["o2"]), "r1-r2", bad(res["r1"], res["r2"]), "o1-r1", bad(res["o1"], res["r1"]), "o2-r2", bad(res["o2"], res["r2"]), "o1-r2", bad(res["o1"], res["r2"]))
