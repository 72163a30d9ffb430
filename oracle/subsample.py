"""TEST INFRASTRUCTURE — sub-sample parity of the full-size BASELINE configurations against the CPU oracle.

The counter-based RNG makes every triangle / ray / pixel independent of the others, so the oracle can redo every k-th
unit of a job the GPU ran in full and must reproduce it (every 1024-th triangle for C3, every 1024-th
ray for C4 with bit-exact ids, every 4096-th pixel for the cameras).  The same run is the timed CPU baseline of the
configuration (SURVEY.md 8(d) "CPU baseline": the extrapolated rate alongside the sample size).

Used by tests/test_gpu_configs.py and by bench.py's cpu_baseline legs only; the product never imports this.
Tolerances: ids and LiDAR distances bit-exact; irradiance, direction and colour <= 1e-5 relative (north_star).
"""
import ctypes
import time

import numpy as np

from raytracerfacility_b200 import structs as S

from . import pyoracle as orc

RTOL = 1e-5
THREADS = 16  # the reference arm and every CPU baseline use min(hardware threads, 16): a box-independent core count


def threads():
    return max(1, min(int(orc.lib().orc_hardware_threads()), THREADS))


def _excess(gpu, cpu, rtol=RTOL, atol=0.0):
    """max over elements of |gpu - cpu| / (atol + rtol |cpu|): <= 1 means within tolerance."""
    g, c = np.asarray(gpu, dtype=np.float64), np.asarray(cpu, dtype=np.float64)
    den = atol + rtol * np.abs(c)
    err = np.abs(g - c)
    with np.errstate(divide="ignore", invalid="ignore"):
        q = np.where(err == 0.0, 0.0, err / den)
    return float(np.nanmax(q)) if q.size else 0.0, float(err.max()) if err.size else 0.0


def _visits(st):
    """Per-ray node visits and triangle tests of the deterministic traversal order on the sample (the GPU's counters equal
    the oracle's bit for bit, tests/test_gpu_trace.py): what a ray of this workload costs, next to M1's 10.95 / 2.29."""
    r = max(1, int(st.rays))
    return {"mean_nodes_per_ray": float(st.nodes) / r, "mean_tris_per_ray": float(st.tris) / r}


def build(cfg, morton_bits=63):
    t0 = time.perf_counter()
    osc = orc.OracleScene(cfg.scene, morton_bits)
    osc.set_brdf(cfg.brdf)
    return osc, time.perf_counter() - t0


def illumination(osc, cfg, gpu_out6, stride, seed, push, nthreads=None):
    """gpu_out6: (n_tris, 6) finalised {energy, direction} of the whole job.  Oracle: triangles 0, stride, 2*stride, ..."""
    nthreads = threads() if nthreads is None else nthreads
    t0 = time.perf_counter()
    cpu, st = osc.illum_mesh(cfg.env, cfg.rp, seed=seed, push=push, tri_stride=stride, nthreads=nthreads)
    dt = time.perf_counter() - t0
    gpu = np.asarray(gpu_out6)[::stride]
    assert gpu.shape == cpu.shape, (gpu.shape, cpu.shape)
    e_q, e_abs = _excess(gpu[:, :3], cpu[:, :3])
    d_q, d_abs = _excess(gpu[:, 3:], cpu[:, 3:], atol=1e-5)
    units = len(cpu) * cfg.rp.samples * 2
    return {"ok": bool(e_q <= 1.0 and d_q <= 1.0), "sample": f"every {stride}-th triangle ({len(cpu)} triangles, {units} paths, "
            f"{int(st.rays)} rays)", "energy_max_err_over_tol": e_q, "energy_max_abs_err": e_abs,
            "direction_max_err_over_tol": d_q, "tolerance": "energy <= 1e-5 relative; direction <= 1e-5 relative + 1e-5",
            "cpu_s": dt, "cpu_threads": nthreads, "cpu_triangle_samples_per_s": units / dt, "cpu_Mrays_per_s": st.rays / dt / 1e6,
            "mean_rays_per_path": float(st.rays) / max(1, int(st.paths)), **_visits(st)}


def lidar(osc, cfg, gpu_t, gpu_prim, stride, ray_begin=0, nthreads=None, strided=False):
    """gpu_t / gpu_prim: the scan's arrays from ray `ray_begin` on.  Bit-exact distances and primitive ids.
    strided=True: the arrays already hold rays 0, stride, 2*stride, ... of the WHOLE grid only (the caller cut the sample
    on the device) — ONE oracle call over a sample that does not depend on how the scan was sharded, which is also what
    the CPU rate is timed on (per-block calls of a few hundred rays each time thread start-up, not tracing)."""
    nthreads = threads() if nthreads is None else nthreads
    total = int(orc.lib().orc_scan_grid_count(ctypes.byref(cfg.grid)))
    if strided:
        assert ray_begin == 0
        t0 = time.perf_counter()
        t, prim, inst, st = osc.scan_grid(cfg.grid, ray_begin=0, ray_end=total, stride=stride, nthreads=nthreads)
        dt = time.perf_counter() - t0
        gt, gp = np.asarray(gpu_t).view(np.uint32), np.asarray(gpu_prim).view(np.uint32)
        assert len(gt) == len(t), (len(gt), len(t))
        same_t = bool(np.array_equal(gt, t.view(np.uint32)))
        same_p = bool(np.array_equal(gp, prim.view(np.uint32)))
        return {"ok": same_t and same_p, "sample": f"every {stride}-th ray of the whole grid ({len(t)} rays, one oracle call)",
                "prim_ids_bit_exact": same_p, "distances_bit_exact": same_t,
                "hit_fraction_on_sample": float((prim != 0xFFFFFFFF).mean()) if len(prim) else 0.0, "cpu_s": dt,
                "cpu_threads": nthreads, "cpu_Mrays_per_s": len(t) / dt / 1e6, **_visits(st)}
    n = len(gpu_t)
    first = (-ray_begin) % stride  # the oracle takes rays 0, stride, ... of the whole grid
    t0 = time.perf_counter()
    t, prim, inst, st = osc.scan_grid(cfg.grid, ray_begin=ray_begin + first, ray_end=min(total, ray_begin + n), stride=stride,
                                      nthreads=nthreads)
    dt = time.perf_counter() - t0
    gt = np.asarray(gpu_t)[first::stride].view(np.uint32)
    gp = np.asarray(gpu_prim)[first::stride].view(np.uint32)
    same_t = bool(np.array_equal(gt, t.view(np.uint32)))
    same_p = bool(np.array_equal(gp, prim.view(np.uint32)))
    return {"ok": same_t and same_p, "sample": f"every {stride}-th ray ({len(t)} rays)", "prim_ids_bit_exact": same_p,
            "distances_bit_exact": same_t, "hit_fraction_on_sample": float((prim != 0xFFFFFFFF).mean()) if len(prim) else 0.0,
            "cpu_s": dt, "cpu_threads": nthreads, "cpu_Mrays_per_s": len(t) / dt / 1e6, **_visits(st)}


def lidar_blocks(osc, cfg, gpu_t, gpu_prim, block_starts, block, stride, nthreads=None):
    """lidar() over interleaved shards: gpu_t / gpu_prim hold the rays of the blocks starting at `block_starts`
    (`block` rays each, the last one possibly shorter), one after the other."""
    parts, off = [], 0
    for b in block_starts:
        n = min(block, len(gpu_t) - off)
        if n <= 0:
            break
        parts.append(lidar(osc, cfg, gpu_t[off:off + n], gpu_prim[off:off + n], stride, ray_begin=b, nthreads=nthreads))
        off += n
    rays = sum(int(p["sample"].split("(")[1].split()[0]) for p in parts)
    cpu_s = sum(p["cpu_s"] for p in parts)
    w = [int(p["sample"].split("(")[1].split()[0]) for p in parts]
    tot = max(1, sum(w))
    return {"ok": all(p["ok"] for p in parts), "sample": f"every {stride}-th ray of {len(parts)} interleaved blocks ({rays} rays)",
            "prim_ids_bit_exact": all(p["prim_ids_bit_exact"] for p in parts),
            "distances_bit_exact": all(p["distances_bit_exact"] for p in parts),
            "hit_fraction_on_sample": sum(p["hit_fraction_on_sample"] * x for p, x in zip(parts, w)) / tot,
            "cpu_s": cpu_s, "cpu_threads": parts[0]["cpu_threads"] if parts else 0, "cpu_Mrays_per_s": rays / max(cpu_s, 1e-12) / 1e6,
            "mean_nodes_per_ray": sum(p["mean_nodes_per_ray"] * x for p, x in zip(parts, w)) / tot,
            "mean_tris_per_ray": sum(p["mean_tris_per_ray"] * x for p, x in zip(parts, w)) / tot}


def camera(osc, cfg, cam_props, gpu_color, stride, nthreads=None):
    """gpu_color: (h, w, 4) colour buffer of frame 0.  Oracle: pixels 0, stride, ... (row-major)."""
    nthreads = threads() if nthreads is None else nthreads
    cp = S.CameraProps.from_buffer_copy(cam_props)
    cp.frame_id = 0
    t0 = time.perf_counter()
    col, _, _, st = osc.camera(cfg.env, cfg.rp, cp, stride=stride, nthreads=nthreads)
    dt = time.perf_counter() - t0
    gpu = np.asarray(gpu_color).reshape(-1, 4)[::stride]
    q, a = _excess(gpu, col)
    paths = len(col) * cfg.rp.samples
    return {"ok": bool(q <= 1.0), "sample": f"every {stride}-th pixel ({len(col)} pixels, {paths} paths, {int(st.rays)} rays)",
            "colour_max_err_over_tol": q, "colour_max_abs_err": a, "tolerance": "<= 1e-5 relative", "cpu_s": dt,
            "cpu_threads": nthreads, "cpu_Mpaths_per_s": paths / dt / 1e6, "cpu_Mrays_per_s": st.rays / dt / 1e6, **_visits(st)}
