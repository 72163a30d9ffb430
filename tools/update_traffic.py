#!/usr/bin/env python
"""Writes profiles/traffic.json from the raw page of an `ncu --set full` capture of k_trace_rays (tools/ncu_trace.sh):
    python tools/update_traffic.py gpurun_out/<tag>_trace_raw.csv profiles/<name>_trace_rays_ncu_raw.csv
The capture is copied to its tracked name under profiles/, and the JSON carries the hash of the kernel sources in THIS
tree (bench.kernel_source_hash): run it on the tree the capture was taken from, before editing a kernel again.  bench.py
drops `roofline.traffic` when the hash no longer matches."""
import csv
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    raw, dst = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(open(raw)))
    hdr, vals = rows[0], rows[2]
    ix = {h: i for i, h in enumerate(hdr)}
    units = rows[1]

    def gb(name):
        v, u = float(vals[ix[name]]), units[ix[name]]
        return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[u]

    rd, wr = gb("dram__bytes_read.sum"), gb("dram__bytes_write.sum")
    if os.path.abspath(raw) != os.path.abspath(dst):
        shutil.copyfile(raw, dst)
    out = {"k_trace_rays_bytes_per_launch": int(rd + wr), "dram_read_bytes": int(rd), "dram_write_bytes": int(wr),
           "log2_rays": 26, "plants": 50, "source_hash": bench.kernel_source_hash(),
           "capture": f"{os.path.relpath(dst, ROOT)}: dram__bytes_read.sum {rd / 1e9:.3f} GB + dram__bytes_write.sum "
                      f"{wr / 1e9:.3f} GB, one launch of 2^26 rays on the C2 scene (tools/quick_trace.py), ncu --set full "
                      f"--clock-control none, cold caches"}
    json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    print(out)


if __name__ == "__main__":
    main()
