"""Times dbk_edge_plane_kernel (svtf_deblock_prepare) on 12 resident 4K mode-info grids, CUDA events per launch."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import harness as H
S = H.S
g = S.make_geom(3840, 2160, 10)
mi = H.pack_mi(H.synth_mode_info(g, 0xB200C3, p_split=0.5, p_skip=0.3, p_inter=0.7))
st = torch.cuda.Stream()
ctxs = []
for i in range(12):
    c = S.FilterContext(g); c.begin(g); c.set_mi(mi); c.sync(); c.set_stream(st.cuda_stream); ctxs.append(c)
def run():
    ev = []
    for c in ctxs:
        e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        e[0].record(st); c.deblock_prepare(); e[1].record(st); ev.append(e)
    torch.cuda.synchronize()
    return statistics.mean(a.elapsed_time(b) for a, b in ev) * 1e3
run(); print(f"edge plane kernel {min(run(), run()):.1f} us per 4K picture (events around one launch)")
