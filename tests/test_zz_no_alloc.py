"""rtf_illumination_flux_mesh_device_static: the flux integration for hosts that cannot let the library allocate.

Who that is: an engine that handed the whole GPU to its own memory pool at start-up (cudaMalloc has nothing left and the
ordinary call's grow-on-first-use workspace fails), and a frame loop that cannot take the device-wide synchronisation
that cudaMalloc / cudaFree imply.  The entry point keeps the partial sums of its reduction tree in an array inside the
library image instead, limited to 2^24 triangles per call.

The check: same inputs through the C ABI into this entry point and into rtf_illumination_flux_mesh_device give the same
eight floats, bit for bit (both run the same fixed-shape tree; bit-exact is the bar the other flux tests use).  The array
is shared and keeps the previous call's contents, so the calls run back to back with different sizes: a long range, a
short one-sided range that leaves most of the long call's partial sums in place, then the long range again.

This file sorts last on purpose (pytest -x stops at the first failure; nothing may hide behind it), and it must pass both
on its own in a fresh process and after the rest of the suite, i.e. with the array untouched or already used.
"""
import numpy as np
import pytest
import torch

from .conftest import assert_bit_equal

pytestmark = pytest.mark.gpu

STATIC_MAX_TRIS = 1 << 24


def _static_flux(rt, d_e6, begin, end, front=True, back=True):
    d_out = torch.full((8,), float("nan"), dtype=torch.float32, device="cuda")  # the caller's own output buffer
    torch.cuda.synchronize()
    rt.illumination_flux_mesh_device_static(d_e6.data_ptr(), begin, end, d_out.data_ptr(), front, back)
    torch.cuda.synchronize()  # the context enqueues on its own non-blocking stream: wait for the whole device
    return d_out.cpu().numpy()


def test_static_scratch_flux_matches_allocating_entry_point(rt, mod, scenes):
    sd = scenes.sorghum_scene(2)
    rt.load_scene(sd, 63)
    n = rt.triangle_count
    assert 2048 < n <= STATIC_MAX_TRIS
    rng = np.random.default_rng(20261106)
    e6_a = rng.random((n, 6), dtype=np.float32)
    e6_b = (rng.random((n, 6), dtype=np.float32) * np.float32(37.0)).astype(np.float32)
    d_a, d_b = torch.from_numpy(e6_a).cuda(), torch.from_numpy(e6_b).cuda()

    # the new entry point goes first, so that in a fresh process nothing else has touched the device for it
    got_long = _static_flux(rt, d_a, 0, n)
    got_short = _static_flux(rt, d_b, 777, 777 + 1500, front=True, back=False)  # 2 tiles: the rest of the array is stale
    got_long_again = _static_flux(rt, d_a, 0, n)
    got_empty = _static_flux(rt, d_a, 5, 5)

    want_long = rt.illumination_flux_mesh_device(d_a.data_ptr(), 0, n)
    want_short = rt.illumination_flux_mesh_device(d_b.data_ptr(), 777, 777 + 1500, front=True, back=False)
    want_empty = rt.illumination_flux_mesh_device(d_a.data_ptr(), 5, 5)
    assert want_long[3] > 0 and np.all(np.isfinite(want_long))
    assert_bit_equal(got_long, want_long, "static flux, first call")
    assert_bit_equal(got_short, want_short, "static flux, short range after a long one")
    assert_bit_equal(got_long_again, want_long, "static flux, long range again")
    assert_bit_equal(got_empty, want_empty, "static flux, empty range")

    # and once more after the allocating path has run, in the other order
    assert_bit_equal(_static_flux(rt, d_b, 0, n), rt.illumination_flux_mesh_device(d_b.data_ptr(), 0, n), "static flux, set b")


def test_static_scratch_flux_refuses_what_it_cannot_hold(rt, mod, scenes):
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, 63)
    d_e6 = torch.zeros((rt.triangle_count, 6), dtype=torch.float32, device="cuda")
    d_out = torch.zeros(8, dtype=torch.float32, device="cuda")
    with pytest.raises(mod.RtfError) as ei:
        rt.illumination_flux_mesh_device_static(d_e6.data_ptr(), 0, STATIC_MAX_TRIS + 1, d_out.data_ptr())
    assert ei.value.code == mod.ERR_LIMIT and "16777216" in str(ei.value)
    with pytest.raises(mod.RtfError) as ei:  # within the limit but past the scene
        rt.illumination_flux_mesh_device_static(d_e6.data_ptr(), 0, rt.triangle_count + 1, d_out.data_ptr())
    assert ei.value.code == mod.ERR_INVALID
    with pytest.raises(mod.RtfError) as ei:
        rt.illumination_flux_mesh_device_static(d_e6.data_ptr(), 0, 1, 0)
    assert ei.value.code == mod.ERR_INVALID
