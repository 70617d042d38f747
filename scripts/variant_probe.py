"""Times the flow sweeps of several tuning builds (build/lib_*.so, scripts/build_variant.sh) in ONE process and checks
every build's output bit for bit against the in-tree library's.  usage: variant_probe.py [size] [lib ...]"""
import ctypes as C
import glob
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import soillib_b200 as soil
from soillib_b200 import silt

args = [a for a in sys.argv[1:]]
n = int(args[0]) if args and args[0].isdigit() else 8192
libs = [a for a in args if a.endswith(".so")] or sorted(glob.glob(os.path.join(ROOT, "build", "lib_*.so")))
libs = [os.path.join(ROOT, "soillib_b200", "libsoilb200.so")] + libs
reps = 30

p = soil.noise_t(); p.seed, p.ext = 3.0, (float(n), float(n))
h = soil.noise(silt.shape(n, n), p, host=silt.gpu).data
small = [torch.rand((a, b), device="cuda") for a, b in ((37, 64), (1000, 1028), (255, 132), (4099, 4096))]
small[0][3, 5] = float("nan"); small[1][0, 0] = 3e38; small[1][7, 9] = -3e38
scale = (C.c_float * 3)(20.0 / n, 20.0 / n, 4.0)
scale2 = (C.c_float * 2)(20.0 / n, 20.0 / n)
GRAD = "--grad" in sys.argv
OPS = (("gradient", 12, 2), ("negslope", 8, 1), ("laplacian", 8, 1)) if GRAD else (("direction", 8, 0), ("steepest", 8, 0), ("normal", 16, 3))


def alloc(t, op):
    ch = dict((o, c) for o, _, c in OPS)[op]
    return torch.empty(tuple(t.shape) + ((ch,) if ch > 1 else ()), device="cuda", dtype=torch.int32 if ch == 0 else torch.float32)


def bind(path):
    lib = C.CDLL(path)
    for f in ("soilb200_direction_f32", "soilb200_steepest_f32"):
        getattr(lib, f).argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    lib.soilb200_normal_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    for f in ("soilb200_gradient_f32", "soilb200_negslope_f32"):
        getattr(lib, f).argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.soilb200_laplacian_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.soilb200_random_weighted_ws.restype = C.c_size_t
    lib.soilb200_random_weighted_ws.argtypes = [C.c_int, C.c_int]
    lib.soilb200_random_weighted_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint64,
                                                 C.c_float, C.c_void_p, C.c_size_t, C.c_void_p]
    return lib


def run(lib, op, t, out):
    a, b = t.shape
    if op == "normal":
        rc = lib.soilb200_normal_f32(t.data_ptr(), out.data_ptr(), a, b, scale, None)
    elif op in ("gradient", "negslope"):
        rc = getattr(lib, f"soilb200_{op}_f32")(t.data_ptr(), out.data_ptr(), a, b, scale2, None)
    elif op == "laplacian":
        rc = lib.soilb200_laplacian_f32(t.data_ptr(), out.data_ptr(), a, b, 1, scale2, None)
    else:
        rc = getattr(lib, f"soilb200_{op}_f32")(t.data_ptr(), out.data_ptr(), a, b, 1, None)
    assert rc == 0, (op, rc)


def timeit(fn):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def outputs(lib, t):
    res = []
    for op, _, _ in OPS:
        o = alloc(t, op)
        run(lib, op, t, o)
        res.append(o)
    torch.cuda.synchronize()
    return res


want = None
for path in libs:
    lib = bind(path)
    got = [outputs(lib, t) for t in [h] + small]
    if want is None:
        want = got
        same = "reference"
    else:
        # bit compare (NaN-safe): view floats as int32
        same = all(torch.equal(a.view(torch.int32), b.view(torch.int32)) for ga, gb in zip(got, want) for a, b in zip(ga, gb))
        same = "bit-identical" if same else "MISMATCH"
    line = f"{os.path.basename(path):24s} {same:14s}"
    for op, nbytes, _ in OPS:
        o = alloc(h, op)
        ms = timeit(lambda: run(lib, op, h, o))
        line += f" {op} {ms:.4f} ms {nbytes * n * n / ms / 1e6:7.1f} GB/s |"
    if "--rw" in sys.argv:
        wsb = lib.soilb200_random_weighted_ws(n, n)
        ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
        o = torch.empty((n, n), device="cuda", dtype=torch.int32)
        k = [0]
        def rw():
            rc = lib.soilb200_random_weighted_f32(h.data_ptr(), o.data_ptr(), n, n, 1, 0, k[0], 10.0, ws.data_ptr(), wsb, None)
            assert rc == 0
            k[0] += 1
        ms = timeit(rw)
        line += f" random_weighted {ms:.4f} ms"
        del ws
    print(line, flush=True)
    del got
