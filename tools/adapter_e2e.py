"""bench.py's adapter end-to-end leg alone (4K 10-bit, 12 pictures per pass); knobs: SVTF_BENCH_ADAPTER_THREADS, SVTF_BENCH_GATHER_THREADS."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import harness as H
import bench
S = H.S
g = S.make_geom(3840, 2160, 10)
wls = [bench.workload(H, g, 0xB200C2 + 7 * i) for i in range(3)]
def barrier(): torch.cuda.synchronize()
r = bench.adapter_e2e_leg(H, S, g, wls, 3, int(os.environ.get("E2E_B", "12")), 5, barrier, 0, np)
print(os.environ.get("SVTF_BENCH_ADAPTER_THREADS", "4"), os.environ.get("SVTF_BENCH_GATHER_THREADS", "8"), r["ms_per_picture"], "ms/picture",
      round(3840 * 2160 / r["ms_per_picture"] / 1e3), "Mpixel/s", r["matches_device_path"])
