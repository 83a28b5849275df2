"""Times the two chain kernels of the 4K 10-bit workload alone (CUDA events, 12 distinct pictures > L2, serialised) and the 4-stream
resident rate, without the rest of bench.py.  Prints one line."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import harness as H
S = H.S
g = S.make_geom(3840, 2160, 10)
wls = []
for i in range(3):
    seed = 0xB200C2 + 7 * i
    rec = H.synth_picture(g, seed)
    mi = H.pack_mi(H.synth_mode_info(g, seed + 1, p_split=0.5, p_skip=0.3, p_inter=0.7))
    cdef, idx = H.synth_cdef(g, seed + 2, n_strengths=8, damping=5)
    ccso, flags = H.synth_ccso(g, seed + 3, quant_idx=(0, 0, 0), support=(0, 1, 2), edge_clf=(0, 0, 0), band_log2=(2, 2, 2), p_on=0.8)
    wls.append((rec, mi, S.lf_params((24, 24), 12, 12), cdef, idx, ccso, flags))
B = 12
NS = int(os.environ.get("NS", "4"))
streams = [torch.cuda.Stream() for _ in range(NS)]
ctxs = []
for i in range(B):
    c = S.FilterContext(g); c.begin(g); w = wls[i % 3]; c.upload(w[0]); c.set_mi(w[1]); c.sync(); ctxs.append(c)
st = streams[0]
def serial():
    ev = []
    for i, c in enumerate(ctxs):
        w = wls[i % 3]; c.set_stream(st.cuda_stream); c.rewind()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record(st); c.loop_filter_frame(w[2]); e[1].record(st); c.filter_chain(None, w[3], w[4], w[5], w[6]); e[2].record(st)
        ev.append(e)
    torch.cuda.synchronize()
    return [statistics.mean(e[k].elapsed_time(e[k + 1]) for e in ev) * 1e3 for k in range(2)]
serial(); a = serial(); b = serial()
for i, c in enumerate(ctxs):
    c.set_stream(streams[i % NS].cuda_stream)
def step():
    for i, c in enumerate(ctxs):
        w = wls[i % 3]; c.rewind(); c.filter_chain(w[2], w[3], w[4], w[5], w[6])
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for s in streams[1:]: s.wait_event(e0)
K = 20
for _ in range(K): step()
for s in streams[1:]:
    e = torch.cuda.Event(); e.record(s); st.wait_event(e)
e1.record(st); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print(f"deblock {min(a[0], b[0]):.1f} us, cdef+ccso {min(a[1], b[1]):.1f} us alone; {NS} streams: {ms / K / B * 1e3:.1f} us per picture, {B * 3840 * 2160 * K / ms / 1e3:.0f} Mpixel/s")
