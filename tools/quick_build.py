#!/usr/bin/env python
"""Developer loop: steady-state rebuild / refit timings of the LBVH builder on the sorghum field.
    python tools/quick_build.py [--plants 50] [--reps 3]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--plants", type=int, default=50)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    import torch

    from raytracerfacility_b200 import module, scenes, structs as S
    rt = module.RayTracer(0)
    sd = scenes.sorghum_scene(args.plants)
    rt.load_scene(sd, 63)
    print("first build", rt.build_timings(), flush=True)
    ih = max(sd.instances)
    g, m, x, c = sd.instances[ih]
    xm = x.copy()
    xm[12] += 0.125
    for refit in (0, 1):
        rt.set_option("refit", refit)
        for rep in range(args.reps):
            rt.set_instance(ih, g, m, xm if rep % 2 == 0 else x, c)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            rt.build_ias(63)
            torch.cuda.synchronize()
            print("refit" if refit else "rebuild", f"wall {1e3 * (time.perf_counter() - t0):.3f} ms", rt.build_timings(), flush=True)
    n_refs = len(rt.debug(S.DBG_REF_TRI))
    print("triangles", rt.triangle_count, "references", n_refs, "nodes", len(rt.debug(S.DBG_WIDE_NODES)))
    rt.destroy()


if __name__ == "__main__":
    main()
