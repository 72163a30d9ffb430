"""bench.py's output contract, on the arm that runs without a GPU: `--impl reference` times the CPU port (the reference has
no CPU implementation of its own and its OptiX build is unavailable on this box) and must print exactly ONE JSON line on
stdout — everything else (library banners, warnings) belongs on stderr — carrying the native arm's metric, unit and config."""
import json
import os
import subprocess
import sys

from .conftest import ROOT

REQUIRED = {"impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
            "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e", "gpu_launches"}


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, out.stdout[:2000]
    d = json.loads(lines[0])
    assert REQUIRED <= set(d), REQUIRED - set(d)
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["gpu_launches"] == 0
    assert d["value"] > 0 and d["steps"] == 1 and d["vs_baseline"] is None
    cfg = d["config"]
    assert cfg["triangles"] == 1000002 and cfg["rays_per_gpu_per_step"] == 1 << 26 and "workload" in cfg
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_native_arm_refuses_to_run_without_a_gpu():
    """No CPU fallback in the product path: without a CUDA device the native arm stops with a message, it does not print a
    number.  (Skipped where a GPU is present.)"""
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0", "--no-extras"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode != 0
    assert not [l for l in out.stdout.splitlines() if l.strip().startswith("{")], out.stdout[:500]
    assert "CUDA" in out.stderr or "cuda" in out.stderr
