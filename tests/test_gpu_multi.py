"""Multi-GPU inside the library (csrc/multi.cu): a C host through rtf_create_multi, the ctypes mirror of the same calls, and
the one-process-per-GPU form (rtf_comm_init_rank under torchrun).  With one visible GPU the single-rank paths still run."""
import os
import subprocess
import sys

import numpy as np
import pytest

from .conftest import ROOT

pytestmark = pytest.mark.gpu


def n_gpus():
    import torch
    return min(torch.cuda.device_count(), 8)


def test_c_host_drives_all_gpus(tmp_path):
    """tests/c/multi_gpu.c: no Python between the host and the GPUs; N-GPU results against one device doing the whole job
    (sample-sharded illumination <= 1e-5, everything else bit-identical)."""
    exe = str(tmp_path / "multi_gpu")
    libdir = os.path.join(ROOT, "raytracerfacility_b200")
    subprocess.check_call(["gcc", "-std=c99", "-O1", "-I", os.path.join(ROOT, "include"), "-o", exe,
                           os.path.join(ROOT, "tests", "c", "multi_gpu.c"), "-L", libdir, "-l:librtf_b200.so", "-lm",
                           "-Wl,-rpath," + libdir])
    n = n_gpus()
    for k in sorted({1, n}):
        out = subprocess.run([exe, str(k)], capture_output=True, text=True, timeout=600)
        assert out.returncode == 0, out.stdout + out.stderr
        assert f"OK: {k} device(s)" in out.stdout


def test_multi_ray_tracer_mirror(mod, S, scenes, orc):
    """module.MultiRayTracer (rtf_create_multi) against the CPU oracle on a small scene."""
    n = n_gpus()
    sd = scenes.sorghum_scene(1)
    mrt = mod.MultiRayTracer(list(range(n)))
    mrt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=8)
    cpu, _ = osc.illum_mesh(env, rp, seed=11, push=1e-3)
    for mode in (0, 1):
        gpu = mrt.estimate_illumination_mesh(env, rp, 11, mode=mode)
        err = np.abs(gpu[:, :3] - cpu[:, :3]) / np.maximum(np.abs(cpu[:, :3]), 1e-30)
        assert err.max() <= 1e-5, (mode, err.max())
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(50000, lo, hi, seed=9)
    assert np.array_equal(mrt.trace_closest(rays).view(np.uint32), osc.trace(rays).view(np.uint32))
    grid = scenes.lidar_grid(center=(0.0, 5.0, 0.0), size=(4.0, 4.0), distance=(0.02, 0.02))
    t, prim, inst = mrt.scan_grid(grid)
    ot, oprim = osc.scan_grid(grid)[:2]
    assert np.array_equal(prim, oprim) and np.array_equal(t.view(np.uint32), ot.view(np.uint32))
    mrt.destroy()


def test_one_process_per_gpu_sharded(tmp_path):
    """torchrun, one rank per GPU: rtf_comm_init_rank + the rtf_*_sharded entry points; rank 0 compares with a plain
    single-GPU context (tests/dist_sharded_check.py)."""
    n = n_gpus()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tests", "dist_sharded_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "SHARDED OK" in out.stdout


_ONE_RANK_FAILS = r"""
import numpy as np
from raytracerfacility_b200 import module as mod, scenes, structs as S
sd = scenes.sorghum_scene(1)
mrt = mod.MultiRayTracer([0, 1])
mrt.ctx(0).load_scene(sd, 63)          # device 1 never commits: its part of every job fails before the collective
env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=8)
for mode in (0, 1):
    try:
        mrt.estimate_illumination_mesh(env, rp, 11, mode=mode)
        raise SystemExit("no error raised")
    except mod.RtfError as e:
        assert e.code == mod.ERR_NOT_READY and "device 1" in str(e), str(e)
cam = mrt.create_camera()
cam.resize(64, 48)
cam.set((0.0, 1.0, 4.0))
assert mrt.render_to_camera(env, cam, rp, outputs_mask=7) is False   # the reference's "no acceleration structure" answer
mrt.ctx(1).load_scene(sd, 63)          # the communicator survived: the same calls now work
a = mrt.estimate_illumination_mesh(env, rp, 11, mode=1)
one = mod.RayTracer(0)
one.load_scene(sd, 63)
b = one.estimate_illumination_mesh(env, rp, seed=11, push=1e-3)
assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
assert mrt.render_to_camera(env, cam, rp, outputs_mask=7) is True
print("AGREEMENT OK")
"""


def test_one_rank_failing_does_not_hang_the_others():
    """A rank that fails before the job's collective (here: device 1 has no committed scene) must make the call return
    an error on every rank instead of leaving the others inside ncclAllReduce / ncclBroadcast.  Runs in a child process
    under a timeout, so a regression shows as a failure, not as a hung suite."""
    if n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    out = subprocess.run([sys.executable, "-c", _ONE_RANK_FAILS], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "AGREEMENT OK" in out.stdout
