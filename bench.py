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
#   always:  R layers 8 + rainfall 4 + uplift 4;
#            W discharge 4 + mass 4 + momentum 8 + debris 4 + debris momentum 8 + layers_next 8 + height 4   = 56
#   albedo:  sediment-free cells copy bedrock colour to the surface (R 12 + W 12)                          = 24
#   per cell a particle touched: read this step's two packed flux records (2 x 16 B; with albedo 2 x 32 B)
# The 16 B/cell gradient/creep record that k_pre hands to the dense kernel is an intermediate of this design, not
# algorithmic: it is counted as traffic (PRE_TRAFFIC_BYTES), not in `achieved`.
DENSE_BYTES, ALBEDO_BYTES, HOT_BYTES, HOT_ALBEDO_BYTES, PRE_TRAFFIC_BYTES = 56, 24, 32, 32, 16
# the whole erosion step, fused ideal (every state array read once and written once,
# all seven flux arrays read and re-zeroed for every cell) and the reference's unfused sequence, bytes per cell-step
STEP_BYTES_IDEAL, STEP_BYTES_IDEAL_NO_ALBEDO, REF_BYTES_PER_CELL = 196, 112, 372


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
    ap.add_argument("--halo", type=int, default=256, help="N>1, peer-mapped: depth of the read-only halo rows (1 = gathers in a neighbour's rows go over NVLink; maxage = 512 keeps them in local HBM, the deeper refresh travels while the interior dense pass runs)")
    ap.add_argument("--no-overlap", action="store_true", help="N>1: exchange the halo rows after the whole dense pass instead of under its interior")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra per-config keys (C1, C2, C5; N>=4: the 8192-row weak-scaling line)")
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
        from soillib_b200.strips import bind_host_to_gpu
        if os.environ.get("SOILB200_BIND"):   # opt-in: the pool's boxes are one NUMA node (nvidia-smi topo -m), no gain there
            bind_host_to_gpu(local)           # before NCCL starts its threads and before any pinned allocation
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
    g = soil.steepest(h, soil.d8)
    decay = silt.tensor.from_torch(torch.full((nx, ny), 0.97, dtype=torch.float32, device="cuda"))
    for name, fn, nbytes in (("direction", lambda: soil.direction(h, soil.d8), 8),
                             ("normal", lambda: soil.normal(h, scale), 16),
                             ("accumulate", lambda: soil.accumulate(g, rain, soil.d8), 12),
                             ("accumulate_decay", lambda: soil.accumulate_decay(g, rain, decay, soil.d8), 16)):
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


def resident_bytes(*objs):
    """Bytes of device memory held by the tensors the listed objects carry (attributes, lists of tensors, silt tensors;
    two levels deep) - what the timed region works on."""
    import torch
    seen, total = set(), 0

    def add(t):
        nonlocal total
        st = t.untyped_storage()
        if t.is_cuda and st.data_ptr() not in seen:
            seen.add(st.data_ptr())
            total += st.nbytes()

    def walk(o, depth):
        if isinstance(o, torch.Tensor):
            add(o)
        elif isinstance(o, (list, tuple)):
            for v in o:
                walk(v, depth)
        elif isinstance(getattr(o, "data", None), torch.Tensor):
            add(o.data)
        elif depth > 0 and hasattr(o, "__dict__") and type(o).__module__.startswith("soillib_b200"):
            for v in vars(o).values():
                walk(v, depth - 1)

    for o in objs:
        walk(o, 2)
    return total


def rooflines(N, albedo, ms_step, dense_ms, particle_ms, hot_cells, traffic=None, traffic_source=None, step_traffic=None,
              step_traffic_source=None):
    """The roofline blocks of the line.  `roofline` is the step by SURVEY.md section 8(d)'s contract figure
    (196 B/cell-step with the colour channels, every flux array read and re-zeroed for every cell); the kernels that make
    up the step follow in `roofline_kernels`."""
    peak, peak_kind = measured_peak()
    ideal = STEP_BYTES_IDEAL if albedo else STEP_BYTES_IDEAL_NO_ALBEDO
    step_bytes = ideal * N
    # what this design has to move per step: dense kernel + k_pre (layers 8 R, record 16 W, re-zeroing of the touched
    # cells' flux records 64 W) with the flux records read for touched cells only
    hot = HOT_BYTES + (HOT_ALBEDO_BYTES if albedo else 0)
    dense_alg = N * (DENSE_BYTES + (ALBEDO_BYTES if albedo else 0)) + hot_cells * hot
    skipped = dense_alg + N * 8 + hot_cells * hot
    gbs = lambda b, ms: b / (ms / 1e3) / 1e9 if ms > 0 else 0.0  # noqa: E731
    step = {"bound": "hbm", "kernel": "whole step: k_transport_both + k_pre + k_fused_step", "achieved": round(gbs(step_bytes, ms_step), 1),
            "peak": peak, "peak_kind": f"of {peak_kind}", "unit": "GB/s", "frac": round(gbs(step_bytes, ms_step) / peak, 4),
            "algorithmic_bytes_per_step": int(step_bytes), "bytes_per_cell_step": ideal,
            "tag_skipped": {"bytes_per_step": int(skipped), "achieved": round(gbs(skipped, ms_step), 1),
                            "frac": round(gbs(skipped, ms_step) / peak, 4),
                            "what": "flux records read and re-zeroed only for the cells a particle touched"},
            "traffic": step_traffic, "traffic_source": step_traffic_source,
            "vs_reference_sequence_bytes": round(gbs(REF_BYTES_PER_CELL * N, ms_step), 1)}
    dense = {"bound": "hbm", "kernel": "k_fused_step", "ms": round(dense_ms, 4), "achieved": round(gbs(dense_alg, dense_ms), 1),
             "peak": peak, "unit": "GB/s", "frac": round(gbs(dense_alg, dense_ms) / peak, 4),
             "algorithmic_bytes_per_launch": int(dense_alg), "hot_cells_per_step": int(hot_cells),
             "intermediate_bytes_per_launch": int(N * PRE_TRAFFIC_BYTES),
             "traffic": traffic, "traffic_source": traffic_source}
    # the particle kernel is bound by the rate of random DRAM requests, not by bytes: ceilings from
    # scripts/micro_gather.cu, per-crossing request counts from the ncu capture (both under profiles/, static)
    part = {"bound": "dram random requests", "kernel": "k_transport_both (+ k_pre under it)", "ms": round(particle_ms, 4),
            "cells_touched_per_step": int(hot_cells),
            "cells_touched_per_s": round(hot_cells / (particle_ms / 1e3) / 1e9, 3) if particle_ms > 0 else None, "unit": "G cells/s",
            "static": {"source": "profiles/r02_transport_ncu_full.txt, profiles/r02_micro_gather.txt (8192^2, 32768 particles)",
                       "crossings_per_launch": 18.7e6, "dram_sectors_read_per_crossing": 7.7, "dram_sectors_written_per_crossing": 1.1,
                       "dram_requests_64B_per_s": 31e9, "dram_random_request_ceiling_per_s": 40e9,
                       "l2_atomic_sectors_per_launch": 77.2e6, "l2_atomic_sectors_per_cycle_per_slice": 0.082,
                       "l2_atomic_peak_sectors_per_cycle_per_slice": 1.0}}
    return step, {"k_fused_step": dense, "k_transport_both": part}


def strips_case(a, rank, world, local, nx_g, ny, P, halo, want_e2e=True):
    """One erosion workload on a (nx_g x ny) map cut into `world` row strips, one per GPU.  Returns a dict of
    measurements on rank 0 (None elsewhere)."""
    import torch
    import soillib_b200 as soil
    from soillib_b200.strips import DistTransport, Strip, erode_strips, partition_rows
    lib = soil._load_library()
    silt = soil.silt
    albedo = not a.no_albedo
    lo, hi = partition_rows(nx_g, world)[rank]
    dev = torch.device("cuda", local)
    halo = 1 if a.no_peers else max(1, min(halo, hi - lo))
    use_view = halo % 2 == 0 and (hi - lo) % 2 == 0 and not a.no_peers and not os.environ.get("SOILB200_NO_VIEW")
    strip = Strip(nx_g, ny, lo, hi, (20.0 / ny, 20.0 / ny, 4.0), P, albedo=albedo, device=dev, halo=halo, view=use_view)
    # every rank synthesises the rows it holds (own + halo rows) of the global fBm DEM: the generator takes the
    # position of a cell, so a block of rows is generated from its first global row
    npar = soil.noise_t()
    npar.seed, npar.ext = 3.0, (float(ny), float(ny))
    full = soil.noise(silt.shape(nx_g, ny), npar, host=silt.gpu).data
    strip.set_from_global(bedrock=full, uplift=full.clamp(0.0, 1.0))
    del full
    torch.cuda.empty_cache()
    p = soil.param_t()
    for k, v in c3_param_values().items():
        setattr(p, k, v)
    tr = strips_case.transport = getattr(strips_case, "transport", None) or DistTransport()
    if not a.no_peers:
        tr.connect([strip])     # particles cross strips through peer-mapped memory (NVLink), no hand-off rounds
    overlap = not a.no_overlap
    N = (hi - lo) * ny
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    erode_strips([strip], tr, p, max(a.warmup, 3), overlap=overlap, export=False)
    barrier(world)
    launches0 = lib.soilb200_launch_count()
    lib.soilb200_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    marks = [] if not a.no_peers else None
    rounds = erode_strips([strip], tr, p, a.steps, events=marks, overlap=overlap, export=False)
    e1.record()
    barrier(world)
    ms_total = max_over_ranks(e0.elapsed_time(e1), world)
    breakdown = None
    if marks:   # per step: particles (+ overlapped precompute) | wait for the other ranks | dense | halos | wait
        seg = [sum(marks[6 * i + j].elapsed_time(marks[6 * i + j + 1]) for i in range(a.steps)) / a.steps for j in range(5)]
        worst = [max_over_ranks(v, world) for v in seg]
        by_rank = None
        if world > 1:
            import torch.distributed as dist
            mine = torch.tensor([seg[0], seg[2]], dtype=torch.float64, device="cuda")
            every = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(every, mine)
            by_rank = {"particle_kernels": [round(float(t[0]), 4) for t in every], "dense": [round(float(t[1]), 4) for t in every]}
        breakdown = {"rank0": [round(v, 4) for v in seg], "max_over_ranks": [round(v, 4) for v in worst], "by_rank": by_rank,
                     "segments": ["particle kernels", "barrier", "dense (halo rows travel under its interior)" if overlap else "dense",
                                  "wait for the halo exchange" if overlap else "halo exchange", "barrier"]}
    launches = lib.soilb200_launch_count() - launches0
    pm, dm, st, hot = C.c_double(), C.c_double(), C.c_int64(), C.c_uint64()
    lib.soilb200_profile_read(C.byref(pm), C.byref(dm), C.byref(st), C.byref(hot))
    lib.soilb200_profile_enable(0)
    clocks = sampler.stop() if rank == 0 else None
    state_bytes = resident_bytes(strip)

    e2e_ms = None
    if want_e2e:
        # end to end: pinned upload of the strip's DEM, K steps each followed by a 16-byte read of the strip's mass sums
        # (accumulated by the dense pass itself), height download
        strip.export_view()
        host_h = torch.empty((strip.rows, ny), dtype=torch.float32).pin_memory()
        host_h.copy_(strip.height)
        host_out = torch.empty((hi - lo, ny), dtype=torch.float32).pin_memory()
        barrier(world)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        stage = torch.empty((strip.rows, ny), dtype=torch.float32, device=dev)
        stage.copy_(host_h, non_blocking=True)
        strip.edit()
        strip.layers[strip.cur][..., 0] = stage - strip.layers[strip.cur][..., 1]
        host_sums = torch.zeros((a.steps, 2), dtype=torch.float64).pin_memory()
        for i in range(a.steps):
            erode_strips([strip], tr, p, 1, mass_sum=True, overlap=overlap, export=False)
            host_sums[i].copy_(strip.mass_sum, non_blocking=True)
        strip.export_view()
        host_out.copy_(strip.own(strip.height), non_blocking=True)
        f1.record()
        barrier(world)
        e2e_ms = max_over_ranks(f0.elapsed_time(f1), world)
        assert bool(torch.isfinite(host_sums).all()) and bool((host_sums[:, 0] != 0).all())
    rows = strip.rows
    if not a.no_peers:
        tr.disconnect([strip])   # unmap the peers' arrays before anyone frees them
    del strip
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    steps_seen = max(st.value, 1)
    hot_cells = hot.value * 1.0 / steps_seen
    ms_step = ms_total / a.steps
    step_roof, kernels = rooflines(N, albedo, ms_step, dm.value / steps_seen, pm.value / steps_seen, hot_cells)
    step_roof["kernel"] += " (rank 0's strip)"
    return {
        "value": round(world * N * a.steps / (ms_total / 1e3) / 1e6, 2), "ms_per_step": round(ms_step, 4),
        "workload": f"erosion step on one {nx_g}x{ny} fBm DEM (seed-z 3) in {world} row strips of {hi - lo} rows, "
                    f"timeStep 10, maxage 512, {P} particles, creep on, albedo {'on' if albedo else 'off'}",
        "parallelism": (f"{world} row strips; halo rows over NCCL send/recv (layers 1 row; "
                        + (f"the packed particle view - heights, discharge, momenta - {halo} rows)" if use_view
                           else f"height, discharge, momenta {halo} rows)")
                        + (", issued after the strip's boundary bands and overlapped with its interior dense pass" if overlap else "")
                        + "; particles cross strips "
                        + ("through peer-mapped (CUDA IPC) memory over NVLink inside one kernel per rank"
                           if not a.no_peers else "by hand-off records over NCCL send/recv")),
        "handoff_rounds_last_step": rounds, "state_bytes_per_gpu": int(state_bytes),
        "gpu_launches": int(launches),
        "phases_ms": {"particles_and_barrier": round(pm.value / steps_seen, 4), "dense": round(dm.value / steps_seen, 4),
                      "halos_and_barrier": round(ms_step - (pm.value + dm.value) / steps_seen, 4), "of": "rank 0",
                      "breakdown": breakdown},
        "roofline": step_roof, "roofline_kernels": kernels,
        "e2e": None if e2e_ms is None else {
            "value": round(world * N * a.steps / (e2e_ms / 1e3) / 1e6, 2), "unit": "Mcell-steps/s",
            "h2d_bytes_per_step": int(rows * ny * 4 / a.steps + 112), "d2h_bytes_per_step": int(N * 4 / a.steps + 16),
            "what": "per rank: pinned strip upload + K steps, each step's mass sums (from the dense pass) copied to pinned host "
                    "memory (stream-ordered) + height download"},
        "clocks": clocks,
    }


def run_ours_strips(a, rank, world, local):
    """N > 1: weak scaling, the per-GPU work of the single-GPU line - one (N * size) x size DEM in N row strips of `size`
    rows, N * particles particles.  That is what `value` reports (the contract's "weak": per-GPU work fixed as N grows).
    For N >= 4 the line also carries BASELINE.json configs[3] as stated - ONE 32768 x 32768 DEM in N row strips (16 x
    the cells and particles of the single-GPU workload, i.e. 16 / N times the per-GPU work; with the fused step's
    scratch it does not fit 2 GPUs) - under the key `config4_32768`."""
    main = strips_case(a, rank, world, local, a.size * world, a.size, a.particles * world, a.halo)
    c4 = None
    if world >= 4 and a.size == 8192 and not a.no_extra:
        density = a.particles / float(a.size * a.size)
        c4 = strips_case(a, rank, world, local, 32768, 32768, int(round(density * 32768 * 32768)), a.halo)
    if rank != 0:
        return None
    line = {
        "metric": "Mcell-steps/s erosion at 8192^2", "value": main["value"], "unit": "Mcell-steps/s", "n_gpus": world,
        "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": main["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": main["workload"], "parallelism": main["parallelism"],
                   "shape": f"(N x {a.size}) x {a.size}: {a.size} rows and {a.particles} particles per GPU, as at N = 1",
                   "handoff_rounds_last_step": main["handoff_rounds_last_step"],
                   "l2": f"inputs larger than L2 ({main['state_bytes_per_gpu'] / 1e9:.1f} GB resident per GPU)"},
        "gpu_launches": main["gpu_launches"], "phases_ms": main["phases_ms"], "roofline": main["roofline"],
        "roofline_kernels": main["roofline_kernels"], "e2e": main["e2e"], "clocks": main["clocks"],
    }
    if c4 is not None:
        line["config4_32768"] = {k: c4[k] for k in ("value", "ms_per_step", "workload", "phases_ms", "e2e", "state_bytes_per_gpu")}
        line["config4_32768"]["cells_per_gpu_vs_n1"] = round(32768.0 * 32768.0 / world / (a.size * a.size), 3)
        line["config4_32768"]["unit"] = "Mcell-steps/s"
    return line


def extra_configs(soil, a):
    """Driver-visible numbers for the BASELINE.json configurations the headline line is not quoted on (1 GPU)."""
    import torch
    silt = soil.silt
    out = {}

    def timed(fn, reps):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    def dem(n):
        npar = soil.noise_t()
        npar.seed, npar.ext = 3.0, (float(n), float(n))
        return soil.noise(silt.shape(n, n), npar, host=silt.gpu)

    # C1: D8 direction + accumulation at 1024^2 (dem_multiflow.py path)
    n = 1024
    h = dem(n)
    rain = silt.tensor.from_torch(torch.ones((n, n), dtype=torch.float32, device="cuda"))
    g = soil.steepest(h, soil.d8)
    d_ms, s_ms = timed(lambda: soil.direction(h, soil.d8), 20), timed(lambda: soil.steepest(h, soil.d8), 20)
    a_ms = timed(lambda: soil.accumulate(g, rain, soil.d8), 20)
    out["c1_flow_1024"] = {"direction_ms": round(d_ms, 4), "steepest_ms": round(s_ms, 4), "accumulate_ms": round(a_ms, 4),
                           "direction_gbs": round(8 * n * n / d_ms / 1e6, 1), "accumulate_gbs": round(12 * n * n / a_ms / 1e6, 1)}
    # C2: erosion_gpu.py defaults at 2048^2 (timeStep 1000, maxage 256, 8192 particles), no colour channels in that script
    n = 2048
    model = soil.map_t(silt.shape(n, n), (20.0 / n, 20.0 / n, 4.0))
    model.height = dem(n)
    p = soil.param_t()
    for k, v in dict(timeStep=1000.0, maxage=256, lrate=1.0, gravity=9.81, uplift=0.01, rainfall=1.0, evapRate=0.0005,
                     fluvialExponent=0.01, exitSlope=0.025).items():
        setattr(p, k, v)
    p.samples = 8192
    data, track = soil.data_t(model.shape), soil.data_t(model.shape)
    ms = timed(lambda: soil.erode(model, data, track, p, 1), 40)
    out["c2_erosion_2048"] = {"ms_per_step": round(ms, 4), "mcell_steps_per_s": round(n * n / ms / 1e3, 1),
                              "workload": "example/erosion_gpu.py:76-91 parameters, 2048^2 fBm DEM, 8192 particles, creep on, albedo off"}
    del model, data, track
    torch.cuda.empty_cache()
    # C5: normal map + multi-flow accumulation (K = 16 samples, T = 10) at 16384^2
    n = 16384
    h = dem(n)
    rain = silt.tensor.from_torch(torch.ones((n, n), dtype=torch.float32, device="cuda"))
    n_ms = timed(lambda: soil.normal(h, (20.0 / n, 20.0 / n, 4.0)), 5)
    t = {}
    for K in (2, 6):
        soil.multiflow(h, rain, soil.d8, K, 10.0)
        t[K] = timed(lambda: soil.multiflow(h, rain, soil.d8, K, 10.0), 1)
    per = (t[6] - t[2]) / 4
    out["c5_flow_16384"] = {"normal_ms": round(n_ms, 4), "normal_gbs": round(16 * n * n / n_ms / 1e6, 1),
                            "multiflow_ms_per_sample": round(per, 3), "multiflow_k16_ms": round(t[2] - 2 * per + 16 * per, 2),
                            "multiflow_algorithmic_gbs": round(40 * n * n / per / 1e6, 1),
                            "what": "K = 16 from the marginal cost per sample plus the one-off seeding of the per-cell streams"}
    return out


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
    state_bytes = resident_bytes(model, data, track)

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
    host_out.copy_(model.surface.data, non_blocking=True)
    f1.record()
    barrier(world)
    e2e_ms = max_over_ranks(f0.elapsed_time(f1), world)
    assert bool(torch.isfinite(host_sums).all()) and bool((host_sums[:, 0] != 0).all())

    if rank != 0:
        return None
    ms_step = ms_total / a.steps
    value = world * N * a.steps / (ms_total / 1e3) / 1e6
    steps_seen = max(st.value, 1)
    hot_cells = hot.value * 1.0 / steps_seen
    traffic, traffic_source, step_traffic, step_traffic_source = None, None, None, None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            t = json.load(f)
            if t.get("size") == a.size and t.get("albedo") == albedo:
                traffic = t.get("k_fused_step_dram_bytes_per_launch")
                traffic_source = "profiles/roofline_traffic.json (static: one ncu --set full capture of this kernel, " + str(t.get("capture", "round 1")) + ")"
                step_traffic = t.get("step_dram_bytes_per_step")
                if step_traffic:
                    step_traffic_source = "profiles/roofline_traffic.json (static: DRAM bytes of the step's three kernels, one ncu --set full capture each)"
    except Exception:  # noqa: BLE001
        pass
    step_roof, kernels = rooflines(N, albedo, ms_step, dm.value / steps_seen, pm.value / steps_seen, hot_cells, traffic, traffic_source,
                                   step_traffic, step_traffic_source)
    line = {
        "metric": "Mcell-steps/s erosion at 8192^2", "value": round(value, 2), "unit": "Mcell-steps/s",
        "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": round(ms_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(nx, ny, a.particles, albedo), "parallelism": "1 GPU",
                   "l2": f"inputs larger than L2 ({state_bytes / 1e9:.1f} GB resident: 132 B/cell of model state + a second layers "
                         "buffer, two sets of packed flux records, the gradient/creep record, tags, generator states)"},
        "gpu_launches": int(launches),
        "phases_ms": {"particles": round(pm.value / steps_seen, 4), "dense": round(dm.value / steps_seen, 4)},
        "roofline": step_roof, "roofline_kernels": kernels,
        "e2e": {"value": round(world * N * a.steps / (e2e_ms / 1e3) / 1e6, 2), "unit": "Mcell-steps/s",
                "h2d_bytes_per_step": int(N * 4 / a.steps + 112), "d2h_bytes_per_step": int(N * 4 / a.steps + 16),
                "what": "pinned DEM upload + K soil.erode() calls, each step's mass sums copied to pinned host memory "
                        "(stream-ordered) + height download"},
        "clocks": clocks,
    }
    del model, data, track
    torch.cuda.empty_cache()
    try:
        line["flow"] = flow_accum_gbs(soil, min(a.size, 8192), min(a.size, 8192))
    except Exception as exc:  # noqa: BLE001
        line["flow"] = {"error": str(exc)}
    if not a.no_extra:
        try:
            line["configs"] = extra_configs(soil, a)
        except Exception as exc:  # noqa: BLE001
            line["configs"] = {"error": str(exc)}
    if not a.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_port(nx, ny, a.particles, albedo)
    return line


def workload_name(nx, ny, particles, albedo=True):
    return (f"erosion step {nx}x{ny} fBm DEM (seed-z 3), timeStep 10, maxage 512, {particles} particles, creep on, "
            f"albedo {'on' if albedo else 'off'}")


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
    torch.cuda.set_device(0)
    # the DEM comes from the oracle's restatement of the reference's FastNoiseLite settings (bit-identical to the product's
    # generator, tests/test_oracle_cpu.py): nothing of soillib_b200 is imported in this arm
    orc.lib().orc_set_threads(0)
    h = torch.from_numpy(orc.noise(nx, ny, 3.0, ext=(float(nx), float(ny)))).cuda()

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
        "config": {"workload": workload_name(nx, ny, a.particles, True), "parallelism": "1 GPU",
                   "implementation": "the reference's own CUDA kernels in their unfused sequence (oracle/_ref/libsoilref.so)",
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
