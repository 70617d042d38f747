"""Generates the fixtures under tests/golden/ from the REFERENCE's own code (test infrastructure).

    python oracle/make_golden.py fnl          # here: reference's vendored FastNoiseLite (oracle/_ref/libfnl.so)
    python oracle/make_golden.py gpu OUTDIR   # on a GPU box: reference's own CUDA kernels (oracle/_ref/libsoilref.so)
    python oracle/make_golden.py gpu_grad OUTDIR   # likewise: gradient family, slope, blur -> ref_gpu_grad.npz
    python oracle/make_golden.py gpu_path OUTDIR   # likewise: solve_uniform -> ref_gpu_path.npz

`fnl` pins the noise restatements (oracle + product).  `gpu` runs the reference
kernels, compiled unmodified behind oracle/shim, on small seeded inputs and
stores inputs and outputs; tests/test_oracle_cpu.py checks the CPU oracle
against them and tests/test_gpu_parity.py checks the product against them.
Inputs are fBm maps from the reference's noise settings (seed-z 3).
"""
from __future__ import annotations

import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLD = os.path.join(ROOT, "tests", "golden")


def fnl():
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libfnl.so"))
    out = {}
    for nx, ny, seed in ((64, 48, 3.0), (33, 37, 0.0)):
        a = np.zeros((nx, ny), np.float32)
        lib.fnl_noise(a.ctypes.data_as(C.c_void_p), nx, ny, C.c_float(1.0), 8, C.c_float(0.6), C.c_float(2.0),
                      C.c_float(seed), C.c_float(nx), C.c_float(ny))
        out[f"fbm_{nx}x{ny}_z{int(seed)}"] = a
    # a 1024^2 map sampled on a coarse lattice: pins large coordinates cheaply
    a = np.zeros((1024, 1024), np.float32)
    lib.fnl_noise(a.ctypes.data_as(C.c_void_p), 1024, 1024, C.c_float(1.0), 8, C.c_float(0.6), C.c_float(2.0),
                  C.c_float(3.0), C.c_float(1024.0), C.c_float(1024.0))
    out["fbm_1024_z3_stride61"] = a[::61, ::61].copy()
    np.savez_compressed(os.path.join(GOLD, "fnl_noise.npz"), **out)
    print("wrote fnl_noise.npz", {k: v.shape for k, v in out.items()})


def gpu(outdir):
    import torch
    import oracle_api as orc
    import ref_api as ref

    os.makedirs(outdir, exist_ok=True)
    dev = "cuda"
    res = {}
    for tag, (nx, ny) in {"a": (96, 80), "b": (37, 53)}.items():
        h = orc.noise(nx, ny, 3.0)
        hd = torch.from_numpy(h).to(dev)
        res[f"{tag}_h"] = h
        for e, en in ((1, "d8"), (0, "d4")):
            res[f"{tag}_direction_{en}"] = ref.direction(hd, e).cpu().numpy()
            res[f"{tag}_steepest_{en}"] = ref.steepest(hd, e).cpu().numpy()
        for k in range(3):
            res[f"{tag}_rw_d8_k{k}"] = ref.random_weighted(hd, 1, 0, k, 10.0).cpu().numpy()
        res[f"{tag}_rw_d4_k0"] = ref.random_weighted(hd, 0, 7, 0, 2.5).cpu().numpy()
        g = ref.steepest(hd, 1)
        rain = torch.ones_like(hd)
        src = torch.from_numpy((np.abs(h) + 0.25).astype(np.float32)).to(dev)
        decay = torch.full_like(hd, 0.9)
        res[f"{tag}_acc_ones"] = ref.accumulate(g, rain, None, 1).cpu().numpy()
        res[f"{tag}_acc_src"] = ref.accumulate(g, src, None, 1).cpu().numpy()
        res[f"{tag}_acc_decay"] = ref.accumulate(g, src, decay, 1).cpu().numpy()
        g4 = ref.steepest(hd, 0)
        res[f"{tag}_acc_ones_d4"] = ref.accumulate(g4, rain, None, 0).cpu().numpy()
        grw = ref.random_weighted(hd, 1, 0, 1, 10.0)
        res[f"{tag}_acc_rw"] = ref.accumulate(grw, rain, None, 1).cpu().numpy()

    # erosion: whole steps of the reference's unfused kernel sequence, both presets, with and without creep
    for tag, (nx, ny, P, param, creep) in {
        "ex": (96, 80, 512, orc.example_param(), True),
        "df": (64, 64, 256, orc.default_param(maxage=128), True),
        "nc": (37, 53, 128, orc.example_param(), False),
    }.items():
        h = orc.noise(nx, ny, 3.0)
        scale = (20.0 / nx, 20.0 / ny, 4.0)
        host = orc.HostModel(h, scale, P, albedo=True, seed=0, uplift=np.clip(h, 0, 1))
        host.layers[..., 1] = 0.01 * (1.0 + np.sin(np.arange(nx * ny, dtype=np.float32).reshape(nx, ny) * 0.37))
        dm = ref.DeviceModel(host, dev)
        res[f"{tag}_layers0"] = host.layers.copy()
        for step in range(3):
            dm.step(param, creep)
            if step in (0, 2):
                for n in ("layers", "height", "waterHeight", "mass", "momentum", "debris", "debrisVel",
                          "albedoSurface", "waterFlux", "massFlux", "debrisFlux"):
                    res[f"{tag}_s{step}_{n}"] = dm.get(n)
        # run-to-run atomic-order noise of the reference itself (SURVEY.md 8c tolerance rule)
        dm2 = ref.DeviceModel(host, dev)
        for step in range(3):
            dm2.step(param, creep)
        res[f"{tag}_s2_layers_rerun"] = dm2.get("layers")
    torch.cuda.synchronize()
    np.savez_compressed(os.path.join(outdir, "ref_gpu.npz"), **res)
    print("wrote", os.path.join(outdir, "ref_gpu.npz"), len(res), "arrays")


def grad_inputs(tag):
    """the inputs of the gradient-family fixtures (shared with tests/test_oracle_cpu.py)"""
    import oracle_api as orc
    nx, ny = {"a": (96, 80), "b": (37, 53)}[tag]
    h = orc.noise(nx, ny, 3.0)
    hn = h.copy()
    hn[(np.arange(nx * ny).reshape(nx, ny) * 7919) % 97 == 0] = np.nan       # nodata cells
    v2 = np.stack([h, np.cos(7.0 * h)], axis=-1).astype(np.float32)
    scale = (20.0 / nx, 20.0 / ny)
    return h, hn, v2, scale


def gpu_grad(outdir):
    """the reference's gradient / negslope / laplacian / slope / gaussian_blur kernels (stock flags)."""
    import torch
    import ref_api as ref

    os.makedirs(outdir, exist_ok=True)
    res = {}
    for tag in ("a", "b"):
        h, hn, v2, scale = grad_inputs(tag)
        hd, hnd, v2d = (torch.from_numpy(x).cuda() for x in (h, hn, v2))
        v1d = hd.unsqueeze(-1).contiguous()
        res[f"{tag}_gradient"] = ref.gradient(hd, scale).cpu().numpy()
        res[f"{tag}_gradient_nan"] = ref.gradient(hnd, scale).cpu().numpy()
        res[f"{tag}_negslope"] = ref.negslope(hd, scale).cpu().numpy()
        res[f"{tag}_negslope_nan"] = ref.negslope(hnd, scale).cpu().numpy()
        res[f"{tag}_laplacian1"] = ref.laplacian(v1d, scale).cpu().numpy()
        res[f"{tag}_laplacian2"] = ref.laplacian(v2d, scale).cpu().numpy()
        res[f"{tag}_slope_d8"] = ref.slope(hd, ref.steepest(hd, 1), scale).cpu().numpy()
        res[f"{tag}_blur1"] = ref.gaussian_blur(v1d, 2.5).cpu().numpy()
        res[f"{tag}_blur2"] = ref.gaussian_blur(v2d, 6.0).cpu().numpy()
    torch.cuda.synchronize()
    np.savez_compressed(os.path.join(outdir, "ref_gpu_grad.npz"), **res)
    print("wrote", os.path.join(outdir, "ref_gpu_grad.npz"), len(res), "arrays")


def path_problem(nx, ny, k):
    """inputs of the solve_uniform fixtures (shared with the tests): downhill flow of an fBm DEM, mild decay"""
    import oracle_api as orc
    h = orc.noise(nx, ny, 3.0).astype(np.float64)
    gx, gy = np.gradient(h)
    flow = (-np.stack([gx, gy], axis=-1) * 50.0).astype(np.float32)
    flow[np.abs(flow).sum(-1) < 1e-6] = (0.01, 0.0)
    source = np.ones((nx, ny, k), np.float32)
    if k == 2:
        source[..., 1] = (1.0 + np.sin(np.arange(nx * ny).reshape(nx, ny) * 0.1)).astype(np.float32)
    decay = np.full((nx, ny), 0.02, np.float32)
    return flow, source, decay


def gpu_path(outdir):
    """the reference's solve_uniform kernels (stock flags), 4096 samples, seed 11."""
    import torch
    import ref_api as ref

    os.makedirs(outdir, exist_ok=True)
    res = {}
    for tag, (nx, ny, k) in {"a": (48, 64, 1), "b": (48, 64, 2), "c": (37, 53, 1)}.items():
        flow, source, decay = path_problem(nx, ny, k)
        f, s, d = (torch.from_numpy(a).cuda() for a in (flow, source, decay))
        st = torch.empty((4096, 48), dtype=torch.uint8, device="cuda")
        ref.seed(st, 11, 0)
        res[f"{tag}_flux"] = ref.solve_uniform(f, s, d, st, (20.0 / nx, 20.0 / ny), 4096).cpu().numpy()
        res[f"{tag}_rng24"] = st[:, :24].cpu().numpy()
    np.savez_compressed(os.path.join(outdir, "ref_gpu_path.npz"), **res)
    print("wrote", os.path.join(outdir, "ref_gpu_path.npz"), len(res), "arrays")


if __name__ == "__main__":
    if sys.argv[1] == "fnl":
        fnl()
    elif sys.argv[1] == "gpu_grad":
        gpu_grad(sys.argv[2])
    elif sys.argv[1] == "gpu_path":
        gpu_path(sys.argv[2])
    else:
        gpu(sys.argv[2])
