"""The C++23 host header (include/soilb200.hpp): tests/cpp/host_api.cpp is compiled with g++ against the C-ABI library,
run on the GPU, and its results are compared with the Python surface on the same DEM (both sit on the same C ABI, so
they must agree bit for bit)."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def compile_host_api(outdir):
    import soillib_b200
    soillib_b200._load_library()                       # builds libsoilb200.so if needed
    libdir = os.path.join(ROOT, "soillib_b200")
    exe = os.path.join(outdir, "host_api")
    cmd = ["g++", "-std=c++23", "-O2", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(CUDA, "include"), "-o", exe, os.path.join(ROOT, "tests", "cpp", "host_api.cpp"),
           "-L", libdir, "-lsoilb200", "-L", os.path.join(CUDA, "lib64"), "-lcudart", f"-Wl,-rpath,{libdir}"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    assert proc.returncode == 0, proc.stderr
    return exe


@pytest.mark.skipif(shutil.which("g++") is None, reason="g++ not available")
def test_host_header_compiles_as_cxx23(tmp_path):
    assert os.path.exists(compile_host_api(str(tmp_path)))


@pytest.mark.gpu
def test_cpp_host_matches_python_surface(tmp_path):
    import soillib_b200 as soil
    from soillib_b200 import silt
    exe = compile_host_api(str(tmp_path))
    proc = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=300)
    assert proc.returncode == 0 and proc.stdout.strip().endswith("OK"), proc.stdout + proc.stderr
    nx, ny = 96, 128
    h = np.fromfile(tmp_path / "height.f32", np.float32).reshape(nx, ny)
    p = soil.noise_t()
    p.seed, p.ext = 3.0, (float(nx), float(ny))
    assert np.array_equal(soil.noise(silt.shape(nx, ny), p).numpy(), h)
    hd = silt.tensor.from_numpy(h).gpu()
    assert np.array_equal(soil.direction(hd, soil.d8).numpy().ravel(), np.fromfile(tmp_path / "direction.i32", np.int32))
    g = soil.steepest(hd, soil.d8)
    assert np.array_equal(g.numpy().ravel(), np.fromfile(tmp_path / "steepest.i32", np.int32))
    rain = silt.tensor.from_numpy(np.ones((nx, ny), np.float32)).gpu()
    assert np.array_equal(soil.accumulate(g, rain, soil.d8).numpy().ravel(),
                          np.fromfile(tmp_path / "accumulate.f32", np.float32))
    assert np.array_equal(soil.normal(hd, (20.0 / nx, 20.0 / ny, 4.0)).numpy().ravel(),
                          np.fromfile(tmp_path / "normal.f32", np.float32))
