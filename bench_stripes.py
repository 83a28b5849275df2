#!/usr/bin/env python
"""bench_stripes.py — SURVEY 8e / config C4: ONE 8K 10-bit picture split across N GPUs by superblock-row stripes.

Launch like bench.py (`python bench_stripes.py` for N=1, torchrun for N>1).  Strong scaling: the work (B 8K pictures,
each split over all ranks) is fixed as N grows.  Apply chain: no collective (each rank uploads its rows + 8 context
rows).  CDEF strength search: per-rank search of its filter-block rows, ONE all_gather of mse/skip (NCCL), then the
same deterministic pick on every rank.  Prints one JSON line on rank 0.

  value   8K pictures' luma pixels per second through deblock -> CDEF+CCSO, surfaces resident, max over ranks
  e2e     same through host buffers (pinned), each rank copying only its rows
  search  filter blocks per second through search (64 strengths) + all_gather + pick
  parity  at start-up rank r's rows are compared with a whole-picture run of the same library on rank 0's GPU
          (N=1: trivially equal) — the bit-exactness claim of 8e, re-checked on the real multi-GPU run."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W8K, H8K, BD = 7680, 4320, 10


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=4, help="8K pictures resident per run (each split over all ranks)")
    ap.add_argument("--width", type=int, default=W8K)
    ap.add_argument("--height", type=int, default=H8K)
    ap.add_argument("--no-search", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    import numpy as np
    import torch
    import harness as H

    S = H.S
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    g = S.make_geom(args.width, args.height, BD)
    npx = args.width * args.height
    B = args.batch
    # every rank builds the SAME pictures (it uploads its rows of them)
    rec = H.synth_picture(g, 0xB200C4)
    soa = H.synth_mode_info(g, 0xB200C4 + 1)
    mi = H.pack_mi(soa)
    lf = S.lf_params((24, 24), 12, 12)
    cdef, idx = H.synth_cdef(g, 0xB200C4 + 2, n_strengths=8, damping=5)
    ccso, flags = H.synth_ccso(g, 0xB200C4 + 3, p_on=0.8)
    parts = S.stripes.partition(S.stripes.sb_rows_of(g), world)
    first, count = parts[rank]
    sgeom = S.stripes.stripe_geom(g, first, count)

    stream = torch.cuda.Stream(device=dev)
    ctxs = []
    for i in range(B):
        c = S.FilterContext(sgeom, device=local)
        c.begin_stripe(g, first, count)
        c.upload(rec)
        c.set_mi(mi)
        c.sync()
        ctxs.append(c)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- parity: striped rows == whole-picture run (same library, rank 0's GPU computes the whole picture) ----
    out = [np.zeros_like(p) for p in rec]
    ctxs[0].filter_chain(lf, cdef, idx, ccso, flags)
    ctxs[0].download(out)
    ctxs[0].rewind()
    r0, r1 = 64 * first, min(g.mi_rows * 4, 64 * (first + count))
    if rank == 0:
        with S.FilterContext(g, device=local) as wc:
            wc.begin(g)
            wc.upload(rec)
            wc.set_mi(mi)
            wc.filter_chain(lf, cdef, idx, ccso, flags)
            whole = wc.download()
    else:
        whole = [np.zeros_like(p) for p in rec]
    if dist is not None:
        for p in range(3):
            t = torch.from_numpy(whole[p].view(np.uint8)).to(dev)
            dist.broadcast(t, 0)
            whole[p] = t.cpu().numpy().view(np.uint16)
    same = all(np.array_equal(out[p][r0 >> s:r1 >> s], whole[p][r0 >> s:r1 >> s]) for p, s in ((0, 0), (1, 1), (2, 1)))
    flag = torch.tensor([1 if same else 0], device=dev)
    if dist is not None:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    parity_ok = bool(flag.item())
    del whole

    # ---- device-resident chain ----
    for c in ctxs:
        c.set_stream(stream.cuda_stream)

    def step():
        for c in ctxs:
            c.rewind()
            c.filter_chain(lf, cdef, idx, ccso, flags)

    for _ in range(args.warmup):
        step()
    barrier()
    l0 = sum(c.launch_count for c in ctxs)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    launches = sum(c.launch_count for c in ctxs) - l0
    tt = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms = float(tt.item())
    value = B * npx * args.steps / (ms * 1e-3) / 1e6

    # ---- e2e: pinned whole-picture buffers, each rank moves its own rows ----
    for c in ctxs:
        c.set_stream(None)
    pin_in, pin_out = S.PinnedPicture(g), [S.PinnedPicture(g) for _ in range(min(B, 2))]
    for p in range(3):
        pin_in.planes[p][...] = rec[p]
    pin_mi = S.PinnedArray(mi.shape, np.uint32)
    pin_mi.array[...] = mi
    L = S._lib.load()

    def e2e_step():
        for j, c in enumerate(ctxs):
            c.begin_stripe(g, first, count)
            c.upload(pin_in.planes)
            c.set_mi(pin_mi.array)
            c.filter_chain(lf, cdef, idx, ccso, flags)
            if j >= 1:
                k = ctxs[j - 1]
                k._chk(L.svtf_download_frame_async(k._h, *k._planes_args(pin_out[(j - 1) % len(pin_out)].planes)))
        k = ctxs[-1]
        k._chk(L.svtf_download_frame_async(k._h, *k._planes_args(pin_out[(B - 1) % len(pin_out)].planes)))
        for c in ctxs:
            c.sync()

    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    tt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    e2e_val = B * npx * args.steps / float(tt.item()) / 1e6
    rows_in = min(g.mi_rows * 4, r1 + 8) - max(0, r0 - 8)
    bytes_row = args.width * 2 * 3 // 2
    h2d = (rows_in * bytes_row + (sgeom.mi_rows * g.mi_cols * 4)) * B
    d2h = (r1 - r0) * bytes_row * B

    # ---- CDEF strength search: stripe search + all_gather + pick ----
    search = None
    if not args.no_search:
        src = H.synth_source(g, rec, 0xB200C4 + 9)
        f1, f2 = H.cdef_level1_lists()
        cfg = S.cdef_search_cfg(f1, f2, base_q_idx=120, lambda_=3000)
        c = ctxs[0]
        c.begin_stripe(g, first, count)
        c.upload(rec)
        c.upload_source(src)
        c.set_mi(mi)
        c.loop_filter_frame(lf)
        c.sync()
        nfb = ((g.mi_rows + 15) // 16) * ((g.mi_cols + 15) // 16)
        bufs = (np.zeros((2, nfb, 64), np.uint64), np.zeros(nfb, np.uint8), np.zeros((nfb, 64), np.uint8), np.zeros((nfb, 64), np.int32))

        def search_step():
            c.cdef_seg_search(cfg, out=bufs)
            m, s = S.stripes.gather_search(bufs[0], bufs[1], g, parts, rank, dist, dev)
            return c.pick_table(cfg, m, s)

        for _ in range(2):
            pick = search_step()
        barrier()
        t0 = time.perf_counter()
        n_s = max(3, args.steps // 2)
        for _ in range(n_s):
            pick = search_step()
        barrier()
        tt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        # every rank must have picked the same strengths
        sig = torch.tensor([pick[0].cdef_bits, *pick[0].cdef_y_strength, *pick[0].cdef_uv_strength, int(pick[1].astype(np.int64).sum())],
                           dtype=torch.int64, device=dev)
        lo, hi = sig.clone(), sig.clone()
        if dist is not None:
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        search = {"value": round(nfb * n_s / float(tt.item()), 1), "unit": "filter blocks/s", "strengths": 64,
                  "includes": "search kernel + D2H of mse + all_gather + finish_cdef_search on every rank",
                  "ms_per_picture": round(float(tt.item()) / n_s * 1e3, 3), "ranks_agree": bool(torch.equal(lo, hi)),
                  "cdef_bits": int(pick[0].cdef_bits)}

    if rank == 0:
        print(json.dumps({
            "metric": "filtered Mpixel/s (deblock+CDEF+CCSO), one 8K picture split by SB-row stripes", "value": round(value, 1),
            "unit": "Mpixel/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u16", "data": "synthetic",
            "config": {"workload": f"single {args.width}x{args.height} 10-bit 4:2:0 picture, {world} stripe(s) of SB rows {[c for _, c in parts]}",
                       "pictures_per_step": B, "l2": f"{B} pictures x 3 surfaces resident per rank",
                       "sharding": "SB-row stripes, 8 context rows re-uploaded per interior boundary, no collective on the apply path"},
            "e2e": {"value": round(e2e_val, 1), "unit": "Mpixel/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "per": "rank"},
            "gpu_launches": int(launches), "search": search, "parity": {"stripes_equal_whole_picture": parity_ok}}))
    for c in ctxs:
        c.close()
    pin_in.close()
    pin_mi.close()
    for p in pin_out:
        p.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0 if parity_ok else 1


if __name__ == "__main__":
    sys.exit(main())
