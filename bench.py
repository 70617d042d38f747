#!/usr/bin/env python
"""bench.py - whole-step erosion throughput (Mcell-steps/s) on N B200s, plus flow-accumulation GB/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--size S]

Workload (BASELINE.json metric: "Mcell-steps/s erosion at 8192^2"): one erosion step = fluvial + debris
particle transport, normalise, mass transfer, creep, apply, merge on a synthetic
fBm DEM (reference noise settings, seed-z 3) with example/erosion_gpu_multiscale.py's parameters
(timeStep 10, maxage 512, 32768 particles) and the colour (albedo) channels HEAD's kernels always carry.
All state (8.9 GB at 8192^2, far larger than the 126 MB L2) is resident in HBM before the timed region.

`--impl reference` times the reference's OWN CUDA kernels (oracle/_ref/libsoilref.so: its graph.cu and
erosion.cu compiled unmodified for sm_100a behind oracle/shim) in their unfused sequence on the same GPU,
same config: soillib ships no CPU implementation of this path, so its stock
code path IS the GPU one.  Without that library or without a GPU it falls back to the CPU oracle port.
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# Algorithmic bytes of the fused dense kernel, per cell and step:
#   always:  R layers 8 + gradient/creep record from k_pre 16 + rainfall 4 + uplift 4;
#            W discharge 4 + mass 4 + momentum 8 + debris 4 + debris momentum 8 + layers_next 8 + height 4   = 72
#   albedo:  sediment-free cells copy bedrock colour to the surface (R 12 + W 12)                          = 24
#   per cell a particle touched: read this step's two packed flux records (2 x 16 B; with albedo 2 x 32 B)
# (k_pre, which runs under the particle kernel, reads layers 8 B, writes its record 16 B and re-zeroes the flux
#  records of the previous step; it is not part of the kernel the roofline line is about)
DENSE_BYTES, ALBEDO_BYTES, HOT_BYTES, HOT_ALBEDO_BYTES = 72, 24, 32, 32
# reference's unfused sequence, compulsory bytes per cell-step
REF_BYTES_PER_CELL = 372


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=8192)
    ap.add_argument("--particles", type=int, default=32768)
    ap.add_argument("--no-albedo", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--halo", type=int, default=1, help="N>1, peer-mapped: depth of the read-only halo rows (1 = gathers in a neighbour's rows go over NVLink; maxage keeps them local but costs a deeper refresh)")
    ap.add_argument("--no-peers", action="store_true", help="N>1: hand particles off with NCCL send/recv instead of peer-mapped memory")
    ap.add_argument("--ref-device", default="auto", choices=["auto", "cuda", "cpu"])
    return ap.parse_args()


def c3_param_values():
    """example/erosion_gpu_multiscale.py:72-98 restricted to fields soil::param_t still has."""
    return dict(timeStep=10.0, maxage=512, lrate=1.0, gravity=9.81, uplift=0.01, rainfall=1.0, evapRate=0.0005,
                fluvialExponent=0.01, exitSlope=0.025)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


def dist_setup(n):
    import torch
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local if world > 1 else 0)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def barrier(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(ms, world):
    if world == 1:
        return ms
    import torch
    import torch.distributed as dist
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def build_workload(soil, nx, ny, particles, albedo):
    """Synthetic model of the C3 configuration, fully resident on the current GPU."""
    import torch
    silt = soil.silt
    npar = soil.noise_t()
    npar.seed, npar.ext = 3.0, (float(nx), float(ny))
    h = soil.noise(silt.shape(nx, ny), npar, host=silt.gpu)
    model = soil.map_t(silt.shape(nx, ny), (20.0 / nx, 20.0 / ny, 4.0))
    model.height = h
    model.uplift = silt.tensor.from_torch(h.data.clamp(0.0, 1.0))   # erosion_gpu_multiscale.py:55-57
    if albedo:
        model.enable_albedo()
    p = soil.param_t()
    for k, v in c3_param_values().items():
        setattr(p, k, v)
    p.samples = particles
    data, track = soil.data_t(model.shape), soil.data_t(model.shape)
    torch.cuda.synchronize()
    return model, data, track, p, h


def flow_accum_gbs(soil, nx, ny, reps=5):
    """Secondary metric: D8 receiver + accumulation of unit rain on the same DEM, algorithmic GB/s."""
    import torch
    silt = soil.silt
    npar = soil.noise_t()
    npar.seed, npar.ext = 3.0, (float(nx), float(ny))
    h = soil.noise(silt.shape(nx, ny), npar, host=silt.gpu)
    rain = silt.tensor.from_torch(torch.ones((nx, ny), dtype=torch.float32, device="cuda"))
    out = {}
    scale = (20.0 / nx, 20.0 / ny, 4.0)
    for name, fn, nbytes in (("direction", lambda: soil.direction(h, soil.d8), 8),
                             ("normal", lambda: soil.normal(h, scale), 16),
                             ("accumulate", None, 12)):
        if fn is None:
            g = soil.steepest(h, soil.d8)
            fn = lambda: soil.accumulate(g, rain, soil.d8)  # noqa: E731
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        out[name] = {"ms": round(ms, 4), "algorithmic_gbs": round(nbytes * nx * ny / ms / 1e6, 1)}
    # multi-flow loop (dem_multiflow.py:43-49 on the device): marginal cost of one more sample, i.e. without the
    # one-off seeding of the per-cell generator streams; per sample R h 4, W+R graph 8, R src 4, W+R acc 8, RW out 16
    t = {}
    for K in (2, 10):
        soil.multiflow(h, rain, soil.d8, K, 10.0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        soil.multiflow(h, rain, soil.d8, K, 10.0)
        e1.record()
        torch.cuda.synchronize()
        t[K] = e0.elapsed_time(e1)
    ms = (t[10] - t[2]) / 8
    out["multiflow_per_sample"] = {"ms": round(ms, 4), "algorithmic_gbs": round(40 * nx * ny / ms / 1e6, 1),
                                   "seeding_ms": round(t[2] - 2 * ms, 3)}
    return out


def cpu_baseline_port(nx, ny, particles, albedo, target_s=12.0):
    """The CPU oracle's erosion step on all host cores, on a bounded sample of the same workload."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_api as orc
    orc.lib().orc_set_threads(0)
    cores = orc.lib().orc_get_threads()
    sn = min(nx, 2048)
    h = orc.noise(sn, sn, 3.0, ext=(float(sn), float(sn)))
    m = orc.HostModel(h, (20.0 / sn, 20.0 / sn, 4.0), particles, albedo=albedo, seed=0, uplift=np.clip(h, 0, 1))
    op = orc.example_param(**{k: v for k, v in c3_param_values().items()})
    m.step(op, True)  # warm
    t0, steps = time.perf_counter(), 0
    while steps < 2 or (time.perf_counter() - t0 < target_s and steps < 50):
        m.step(op, True)
        steps += 1
    dt = time.perf_counter() - t0
    return {"value": round(sn * sn * steps / dt / 1e6, 3), "unit": "Mcell-steps/s", "cores": int(cores), "kind": "port",
            "sample": f"{steps} steps of the same step composition on a {sn}x{sn} fBm DEM, {particles} particles, "
                      f"oracle/liboracle.so with OpenMP on {cores} threads"}


def run_ours_strips(a, rank, world, local):
    """N > 1: one (N*size) x size map cut into N row strips of `size` rows, one per GPU (weak scaling: the
    per-GPU rows and the particle density stay fixed).  Halo rows and particle hand-off go over NCCL."""
    import torch
    import soillib_b200 as soil
    from soillib_b200.strips import DistTransport, Strip, erode_strips, partition_rows
    lib = soil._load_library()
    silt = soil.silt
    ny, nx_g = a.size, a.size * world
    albedo = not a.no_albedo
    P = a.particles * world
    lo, hi = partition_rows(nx_g, world)[rank]
    dev = torch.device("cuda", local)
    # peer-mapped: read-only halo rows as deep as a particle can travel (one row per crossing), so that every
    # gather stays in local HBM and only the posted deposits cross NVLink
    halo = 1 if a.no_peers else max(1, min(a.halo, a.size))
    strip = Strip(nx_g, ny, lo, hi, (20.0 / ny, 20.0 / ny, 4.0), P, albedo=albedo, device=dev, halo=halo)
    # every rank synthesises its own rows (+ halo rows) of the global fBm DEM
    npar = soil.noise_t()
    npar.seed, npar.ext = 3.0, (float(ny), float(ny))
    full = soil.noise(silt.shape(nx_g, ny), npar, host=silt.gpu).data
    strip.set_from_global(bedrock=full, uplift=full.clamp(0.0, 1.0))
    del full
    torch.cuda.empty_cache()
    p = soil.param_t()
    for k, v in c3_param_values().items():
        setattr(p, k, v)
    tr = DistTransport()
    if not a.no_peers:
        tr.connect([strip])     # particles cross strips through peer-mapped memory (NVLink), no hand-off rounds
    N = (hi - lo) * ny
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    erode_strips([strip], tr, p, max(a.warmup, 3))
    barrier(world)
    launches0 = lib.soilb200_launch_count()
    lib.soilb200_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    marks = [] if not a.no_peers else None
    rounds = erode_strips([strip], tr, p, a.steps, events=marks)
    e1.record()
    barrier(world)
    ms_total = max_over_ranks(e0.elapsed_time(e1), world)
    breakdown = None
    if marks:   # per step: particles (+ overlapped precompute) | wait for the other ranks | dense | halos | wait
        seg = [sum(marks[6 * i + j].elapsed_time(marks[6 * i + j + 1]) for i in range(a.steps)) / a.steps for j in range(5)]
        worst = [max_over_ranks(v, world) for v in seg]
        breakdown = {"rank0": [round(v, 4) for v in seg], "max_over_ranks": [round(v, 4) for v in worst],
                     "segments": ["particle kernels", "barrier", "dense", "halo exchange", "barrier"]}
    launches = lib.soilb200_launch_count() - launches0
    pm, dm, st, hot = C.c_double(), C.c_double(), C.c_int64(), C.c_uint64()
    lib.soilb200_profile_read(C.byref(pm), C.byref(dm), C.byref(st), C.byref(hot))
    lib.soilb200_profile_enable(0)
    clocks = sampler.stop() if rank == 0 else None

    # end to end: pinned upload of the strip's DEM, K steps each followed by a 16-byte mass read, height download
    host_h = torch.empty((strip.rows, ny), dtype=torch.float32).pin_memory()
    host_h.copy_(strip.height)
    host_out = torch.empty((hi - lo, ny), dtype=torch.float32).pin_memory()
    barrier(world)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    stage = torch.empty((strip.rows, ny), dtype=torch.float32, device=dev)
    stage.copy_(host_h, non_blocking=True)
    strip.layers[strip.cur][..., 0] = stage - strip.layers[strip.cur][..., 1]
    host_sums = torch.zeros((a.steps, 2), dtype=torch.float64).pin_memory()
    for i in range(a.steps):
        erode_strips([strip], tr, p, 1)
        L = strip.own(strip.layers[strip.cur])
        host_sums[i].copy_(torch.stack([L[..., 0].double().sum(), L[..., 1].double().sum()]), non_blocking=True)
    host_out.copy_(strip.own(strip.height), non_blocking=True)
    f1.record()
    barrier(world)
    e2e_ms = max_over_ranks(f0.elapsed_time(f1), world)
    assert bool(torch.isfinite(host_sums).all())
    if rank != 0:
        return None
    # rank 0's strip: the dense kernel against the HBM roofline, as at N=1
    steps_seen = max(st.value, 1)
    hot_cells = hot.value * 1.0 / steps_seen
    bytes_dense = N * (DENSE_BYTES + (ALBEDO_BYTES if albedo else 0)) + hot_cells * (HOT_BYTES + (HOT_ALBEDO_BYTES if albedo else 0))
    dense_ms = dm.value / steps_seen
    peak, peak_kind = measured_peak()
    achieved = bytes_dense / (dense_ms / 1e3) / 1e9 if dense_ms > 0 else 0.0
    return {
        "metric": "Mcell-steps/s erosion at 8192^2", "value": round(world * N * a.steps / (ms_total / 1e3) / 1e6, 2),
        "unit": "Mcell-steps/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
        "ms_per_step": round(ms_total / a.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"erosion step on one {nx_g}x{ny} fBm DEM (seed-z 3) in {world} row strips of {a.size} rows, "
                               f"timeStep 10, maxage 512, {P} particles, creep on, albedo {'on' if albedo else 'off'}",
                   "parallelism": (f"{world} row strips; halo rows over NCCL send/recv (layers 1 row; height, discharge, momenta "
                                   f"{halo} rows); particles cross strips "
                                   + ("through peer-mapped (CUDA IPC) memory over NVLink inside one kernel per rank"
                                      if not a.no_peers else "by hand-off records over NCCL send/recv")),
                   "handoff_rounds_last_step": rounds,
                   "l2": f"inputs larger than L2 ({N * 132 / 1e9:.1f} GB of state per GPU)"},
        "gpu_launches": int(launches),
        "phases_ms": {"particles_and_barrier": round(pm.value / steps_seen, 4), "dense": round(dense_ms, 4),
                      "halos_and_barrier": round(ms_total / a.steps - (pm.value + dm.value) / steps_seen, 4),
                      "of": "rank 0", "breakdown": breakdown},
        "roofline": {"bound": "hbm", "kernel": "k_fused_step (rank 0's strip)", "achieved": round(achieved, 1), "peak": peak,
                     "peak_kind": f"of {peak_kind}", "unit": "GB/s", "frac": round(achieved / peak, 4),
                     "algorithmic_bytes_per_launch": int(bytes_dense), "hot_cells_per_step": int(hot_cells),
                     "traffic": None},
        "e2e": {"value": round(world * N * a.steps / (e2e_ms / 1e3) / 1e6, 2), "unit": "Mcell-steps/s",
                "h2d_bytes_per_step": int(strip.rows * ny * 4 / a.steps + 112), "d2h_bytes_per_step": int(N * 4 / a.steps + 16),
                "what": "per rank: pinned strip upload + K steps, each step's mass sums copied to pinned host memory "
                        "(stream-ordered) + height download"},
        "clocks": clocks,
    }


def run_ours(a):
    import torch
    rank, world, local = dist_setup(a.gpus)
    if world > 1:
        return run_ours_strips(a, rank, world, local)
    import soillib_b200 as soil
    lib = soil._load_library()
    nx = ny = a.size
    albedo = not a.no_albedo
    model, data, track, p, h = build_workload(soil, nx, ny, a.particles, albedo)
    N = nx * ny

    # clocks are sampled from the warm-up on: the timed region is ~0.1 s, too short for nvidia-smi alone
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(a.warmup, 3)):
        soil.erode(model, data, track, p, 1)
    barrier(world)
    launches0 = lib.soilb200_launch_count()
    lib.soilb200_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    soil.erode(model, data, track, p, a.steps)
    e1.record()
    barrier(world)
    ms_total = max_over_ranks(e0.elapsed_time(e1), world)
    launches = lib.soilb200_launch_count() - launches0
    pm, dm, st, hot = C.c_double(), C.c_double(), C.c_int64(), C.c_uint64()
    lib.soilb200_profile_read(C.byref(pm), C.byref(dm), C.byref(st), C.byref(hot))
    lib.soilb200_profile_enable(0)
    clocks = sampler.stop() if rank == 0 else None

    # end to end through the public API: upload the DEM from pinned memory, K x soil.erode(...) each
    # followed by a device->host read of the step's mass sums, download the height
    host_h = torch.empty((nx, ny), dtype=torch.float32).pin_memory()
    host_h.copy_(h.data)
    host_out = torch.empty((nx, ny), dtype=torch.float32).pin_memory()
    barrier(world)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    stage = torch.empty((nx, ny), dtype=torch.float32, device="cuda")
    stage.copy_(host_h, non_blocking=True)
    model.height = soil.silt.tensor.from_torch(stage)
    # every step's result (its mass sums, 16 bytes) goes to pinned host memory as a stream-ordered copy: read every
    # step, but the host does not stall the GPU between steps; checked after the timed region
    host_sums = torch.zeros((a.steps, 2), dtype=torch.float64).pin_memory()
    for i in range(a.steps):
        host_sums[i].copy_(soil.erode(model, data, track, p, 1, mass_sum=True), non_blocking=True)
    host_out.copy_(model.height.data, non_blocking=True)
    f1.record()
    barrier(world)
    e2e_ms = max_over_ranks(f0.elapsed_time(f1), world)
    assert bool(torch.isfinite(host_sums).all()) and bool((host_sums[:, 0] != 0).all())

    if rank != 0:
        return None
    ms_step = ms_total / a.steps
    value = world * N * a.steps / (ms_total / 1e3) / 1e6
    hot_cells = hot.value * 1.0 / max(st.value, 1)
    bytes_dense = N * (DENSE_BYTES + (ALBEDO_BYTES if albedo else 0)) + hot_cells * (HOT_BYTES + (HOT_ALBEDO_BYTES if albedo else 0))
    dense_ms = dm.value / max(st.value, 1)
    peak, peak_kind = measured_peak()
    achieved = bytes_dense / (dense_ms / 1e3) / 1e9 if dense_ms > 0 else 0.0
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            t = json.load(f)
            if t.get("size") == a.size and t.get("albedo") == albedo:
                traffic = t.get("k_fused_step_dram_bytes_per_launch")
    except Exception:  # noqa: BLE001
        pass
    line = {
        "metric": "Mcell-steps/s erosion at 8192^2", "value": round(value, 2), "unit": "Mcell-steps/s",
        "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": round(ms_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"erosion step {nx}x{ny} fBm DEM (seed-z 3), timeStep 10, maxage 512, "
                               f"{a.particles} particles, creep on, albedo {'on' if albedo else 'off'}",
                   "parallelism": "1 GPU" if world == 1 else f"{world} independent {nx}x{ny} domains (replicas, no halo exchange)",
                   "l2": f"inputs larger than L2 ({N * 132 / 1e9:.1f} GB of state)"},
        "gpu_launches": int(launches),
        "phases_ms": {"particles": round(pm.value / max(st.value, 1), 4), "dense": round(dense_ms, 4)},
        "roofline": {"bound": "hbm", "kernel": "k_fused_step", "achieved": round(achieved, 1), "peak": peak,
                     "peak_kind": f"of {peak_kind}", "unit": "GB/s", "frac": round(achieved / peak, 4),
                     "algorithmic_bytes_per_launch": int(bytes_dense), "hot_cells_per_step": int(hot_cells),
                     "traffic": traffic,
                     "whole_step_vs_reference_bytes": round(REF_BYTES_PER_CELL * N / (ms_step / 1e3) / 1e9, 1)},
        "e2e": {"value": round(world * N * a.steps / (e2e_ms / 1e3) / 1e6, 2), "unit": "Mcell-steps/s",
                "h2d_bytes_per_step": int(N * 4 / a.steps + 112), "d2h_bytes_per_step": int(N * 4 / a.steps + 16),
                "what": "pinned DEM upload + K soil.erode() calls, each step's mass sums copied to pinned host memory "
                        "(stream-ordered) + height download"},
        "clocks": clocks,
    }
    if world == 1:
        try:
            line["flow"] = flow_accum_gbs(soil, min(a.size, 8192), min(a.size, 8192))
        except Exception as exc:  # noqa: BLE001
            line["flow"] = {"error": str(exc)}
        if not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_port(nx, ny, a.particles, albedo)
    return line


def run_reference(a):
    """The reference's own implementation of the same step on the same config."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return None  # rank 0 alone runs the reference arm
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    nx = ny = a.size
    N = nx * ny
    use_cuda = False
    if a.ref_device in ("auto", "cuda"):
        try:
            import torch
            import ref_api
            use_cuda = torch.cuda.is_available() and ref_api.available()
        except Exception:  # noqa: BLE001
            use_cuda = False
    base = {
        "impl": "reference", "metric": "Mcell-steps/s erosion at 8192^2", "unit": "Mcell-steps/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": max(a.warmup, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
    }
    if not use_cuda:
        cb = cpu_baseline_port(nx, ny, a.particles, True, target_s=20.0)
        base.update({"value": cb["value"], "ms_per_step": None, "cpu_baseline": cb,
                     "config": {"workload": "CPU port of the reference step (reference kernels unavailable on this host): " + cb["sample"]},
                     "e2e": {"value": cb["value"], "unit": "Mcell-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        return base

    import numpy as np
    import torch
    import oracle_api as orc
    import ref_api
    import soillib_b200 as soil  # only for the fBm DEM generator (input synthesis, not the timed path)
    torch.cuda.set_device(0)
    npar = soil.noise_t()
    npar.seed, npar.ext = 3.0, (float(nx), float(ny))
    h = soil.noise(soil.silt.shape(nx, ny), npar, host=soil.silt.gpu).data

    z = lambda *s: torch.zeros(s, dtype=torch.float32, device="cuda")  # noqa: E731
    t = {"layers": z(nx, ny, 2), "height": h.clone(), "rainfall": torch.ones((nx, ny), device="cuda"),
         "uplift": h.clamp(0.0, 1.0), "waterHeight": z(nx, ny), "waterFlux": z(nx, ny), "mass": z(nx, ny),
         "massFlux": z(nx, ny), "momentum": z(nx, ny, 2), "velocityFlux": z(nx, ny, 2), "debris": z(nx, ny),
         "debrisFlux": z(nx, ny), "debrisVel": z(nx, ny, 2), "debrisVelFlux": z(nx, ny, 2),
         "albedoBedrock": torch.full((nx, ny, 3), 0.5, device="cuda"), "albedoSurface": torch.full((nx, ny, 3), 0.5, device="cuda"),
         "albedoFluxFluvial": z(nx, ny, 3), "albedoFluxDebris": z(nx, ny, 3), "delta": z(nx, ny, 2),
         "rngFluvial": torch.from_numpy(orc.seed_states(a.particles, 0)).cuda(),
         "rngDebris": torch.from_numpy(orc.seed_states(a.particles, 1)).cuda()}
    t["layers"][..., 0] = h
    m = ref_api.RefModel()
    m.nx, m.ny, m.P = nx, ny, a.particles
    m.scale[0], m.scale[1], m.scale[2] = 20.0 / nx, 20.0 / ny, 4.0
    for name, _ in ref_api.RefModel._fields_[3:-1]:
        setattr(m, name, t[name].data_ptr())
    op = orc.example_param(**c3_param_values())
    lib = ref_api.lib()
    for _ in range(max(a.warmup, 3)):
        lib.ref_erode_step(C.byref(m), C.byref(op), 1)
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()   # the reference launches on the legacy default stream, which torch's default stream is
    for _ in range(a.steps):
        rc = lib.ref_erode_step(C.byref(m), C.byref(op), 1)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    assert rc == 0 and bool(torch.isfinite(t["layers"]).all())
    ms = e0.elapsed_time(e1)
    value = N * a.steps / (ms / 1e3) / 1e6
    base.update({
        "value": round(value, 2), "ms_per_step": round(ms / a.steps, 4),
        "config": {"workload": f"erosion step {nx}x{ny} fBm DEM (seed-z 3), timeStep 10, maxage 512, {a.particles} "
                               "particles, creep on, albedo on - reference's unfused kernel sequence on this GPU",
                   "device": "cuda:0 (soillib ships no CPU path for erosion/flow; its stock path is CUDA)"},
        "cpu_baseline": {"value": round(value, 2), "unit": "Mcell-steps/s", "cores": 0, "kind": "reference",
                         "sample": "full workload; reference's own CUDA kernels (oracle/_ref/libsoilref.so), not host cores"},
        "e2e": {"value": round(value, 2), "unit": "Mcell-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "roofline": {"bound": "hbm", "achieved": round(REF_BYTES_PER_CELL * N / (ms / a.steps / 1e3) / 1e9, 1),
                     "peak": measured_peak()[0], "unit": "GB/s", "frac": round(REF_BYTES_PER_CELL * N / (ms / a.steps / 1e3) / 1e9 / measured_peak()[0], 4),
                     "traffic": None, "note": "372 B/cell-step compulsory bytes of the unfused sequence (SURVEY.md App. C)"},
        "clocks": clocks,
    })
    return base


def main():
    a = parse()
    line = run_reference(a) if a.impl == "reference" else run_ours(a)
    if line is not None:
        print(json.dumps(line), flush=True)
    if int(os.environ.get("WORLD_SIZE", 1)) > 1 and a.impl == "ours":
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
