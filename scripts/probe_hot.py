"""probe: is the fused kernel's hot (flux read + re-zero) path slow by itself, or only after atomics?"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import soillib_b200 as soil

lib = soil._load_library()
for albedo in (True, False):
    for P, preset in ((32, True), (32, False), (32768, False)):
        model, data, track, p, h = bench.build_workload(soil, 4096, 4096, P, albedo)
        soil.erode(model, data, track, p, 2)
        torch.cuda.synchronize()
        lib.soilb200_profile_enable(1)
        for _ in range(5):
            if preset:
                model._state.tags[model._state.cur].fill_(1)
            soil.erode(model, data, track, p, 1)
        pm, dm, st, hot = C.c_double(), C.c_double(), C.c_int64(), C.c_uint64()
        lib.soilb200_profile_read(C.byref(pm), C.byref(dm), C.byref(st), C.byref(hot))
        lib.soilb200_profile_enable(0)
        print(f"albedo={albedo} P={P} preset_tags={preset}: particles {pm.value/st.value:.3f} ms dense {dm.value/st.value:.3f} ms hot_cells {hot.value/st.value:.0f}", flush=True)
        del model, data, track
        torch.cuda.empty_cache()
