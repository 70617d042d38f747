"""Host-side tensor helpers of the silt stand-in that need no GPU (soillib_b200/silt.py)."""
import numpy as np
import pytest


def test_resize_between_resolutions():
    """soil.resize(target, source, newres, oldres), example/erosion_gpu_multiscale.py:112-138 (definition: silt.py)."""
    import soillib_b200 as soil
    from soillib_b200 import silt

    old, new = (8, 12), (16, 24)
    x, y = np.meshgrid(np.arange(old[0], dtype=np.float32), np.arange(old[1], dtype=np.float32), indexing="ij")
    plane = (2.0 * x + 3.0 * y + 1.0).astype(np.float32)

    # flat buffers, as the script allocates them (soil.buffer(soil.float32, index.elem(), ...))
    src = silt.tensor.from_numpy(plane.reshape(-1).copy())
    dst = soil.buffer(soil.float32, soil.index(*new).elem(), soil.cpu)
    soil.resize(dst, src, new, old)
    got = dst.numpy().reshape(new)
    # a linear field is reproduced away from the edges: the new cell centre (i + 0.5) / 2 - 0.5 in old cell units
    xi = (np.arange(new[0], dtype=np.float64) + 0.5) / 2 - 0.5
    yi = (np.arange(new[1], dtype=np.float64) + 0.5) / 2 - 0.5
    want = 2.0 * xi[:, None] + 3.0 * yi[None, :] + 1.0
    np.testing.assert_allclose(got[1:-1, 1:-1], want[1:-1, 1:-1], rtol=1e-6)
    assert got.min() >= plane.min() and got.max() <= plane.max()          # edges continue, nothing overshoots

    # a constant stays a constant; equal resolutions copy; interleaved components stay with their cell
    const = silt.tensor.from_numpy(np.full(old, 0.25, np.float32))
    up = silt.tensor(silt.float32, silt.shape(*new), silt.cpu)
    silt.resize(up, const, new, old)
    assert np.array_equal(up.numpy(), np.full(new, 0.25, np.float32))
    same = silt.tensor(silt.float32, silt.shape(*old), silt.cpu)
    silt.resize(same, silt.tensor.from_numpy(plane), old, old)
    assert np.array_equal(same.numpy(), plane)
    mom = np.stack([plane, -plane], axis=-1)
    up2 = silt.tensor(silt.float32, silt.shape(*new, 2), silt.cpu)
    silt.resize(up2, silt.tensor.from_numpy(mom), new, old)
    assert np.array_equal(up2.numpy()[..., 0], got) and np.array_equal(up2.numpy()[..., 1], -got)

    with pytest.raises(ValueError):
        silt.resize(up, const, (5, 5), old)


def test_erosion_gpu_script_host_side():
    """examples/erosion_gpu.py (the reference's example/erosion_gpu.py on this library): everything that runs without a
    GPU - the parameter block with the previous API's names, the host tensors - and the error a host without one gets."""
    import os
    import sys
    import torch
    import soillib_b200 as soil
    from soillib_b200 import silt
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples"))
    import erosion_gpu

    p = erosion_gpu.parameters()
    assert p.maxage == 256 and p.timeStep == 1000.0 and abs(p.exitSlope - 0.025) < 1e-9
    assert p.legacy["samples"] == 8192 and "debrisYieldStress" in p.legacy   # recorded, not silently dropped
    t = erosion_gpu.full(1.0, silt.shape(8, 6))
    assert t.numpy().shape == (8, 6) and float(t.numpy().min()) == 1.0
    m = erosion_gpu.full(0.0, silt.shape(8, 6, 2))
    assert m.numpy().shape == (8, 6, 2)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError):
            erosion_gpu.main(res=16, steps=1, quiet=True)


def test_multiscale_script_host_side():
    """examples/erosion_gpu_multiscale.py: level list and parameter block (no GPU needed)."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples"))
    import erosion_gpu_multiscale as ms

    assert ms.default_levels() == [([128, 128], 2048), ([256, 256], 4), ([1000, 1000], 4)]     # the script's list
    lv = ms.default_levels(8192)
    assert [r[0] for r, _ in lv] == [128, 256, 512, 1024, 2048, 4096, 8192]                    # BASELINE configuration 3
    p = ms.parameters()
    assert p.timeStep == 10.0 and p.maxage == 512 and p.legacy["samples"] == 32768
    assert ms.pixel_scale([128, 128]) == [20.0 / 128, 20.0 / 128, 4.0]


def test_data_t_fields_exist_without_gpu_state():
    """soil.data_t: the five fields are properties (allocated at first access on a GPU; None without one)."""
    import torch
    import soillib_b200 as soil
    d = soil.data_t(soil.index(4, 6))
    assert set(soil.data_t._FIELDS) == {"discharge", "mass", "debris", "momentum", "debris_momentum"}
    if not torch.cuda.is_available():
        assert d.discharge is None
    t = soil.silt.tensor.from_numpy(np.zeros((4, 6), np.float32))
    d.mass = t
    assert d.mass is t
