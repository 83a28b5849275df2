"""The bench's CDEF-search leg alone (pictures in flight vs one at a time), for A/B runs of the pick path."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import harness as H
import bench
S = H.S
g = S.make_geom(3840, 2160, 10)
wl = bench.workload(H, g, 0xB200C2)
n = int(os.environ.get("INFLIGHT", "8"))
ctxs = [S.FilterContext(g) for _ in range(n)]
streams = [torch.cuda.Stream() for _ in range(n)]
for c, s in zip(ctxs, streams): c.set_stream(s.cuda_stream)
out = bench.gpu_search_leg(H, S, g, wl, ctxs, 5, np)
print(json.dumps({k: out[k] for k in ("value", "e2e", "one_picture_at_a_time") if k in out}))
