#!/usr/bin/env python
"""bench.py — filtered Mpixel/s of the in-loop filter chain (deblock -> CDEF -> CCSO).

Workload at every N: BASELINE.json configs[1], "decoder post-recon filter chain on 4K 10-bit 4:2:0
synthetic stream": a batch of distinct 3840x2160 10-bit pictures per GPU (kernel-level generator of
SURVEY §8d C2: quadtree mode info, filter_level 24/24/12/12, 8 CDEF strength pairs, CCSO on all
three planes).  One step = one pass of the chain over the batch.

  value     device-resident: pictures and side info already in HBM, chain kernels only, timed with
            CUDA events on the launching stream; batch > L2 (126 MB) so every step streams from HBM.
  e2e       the same metric through the C ABI with HOST buffers: per picture upload (pinned) +
            mi upload + chain + download, pipelined over several contexts/streams; wall clock
            bracketed by device synchronisation.
  roofline  dominant kernel of the chain: algorithmic bytes (deblock 3b+0.5, CDEF 3b,
            CCSO 4b per luma pixel, b = bytes/sample) / its mean launch time (CUDA events inside the
            timed region) against MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline  the reference's own C (oracle/_ref) — or the oracle port if that library is absent —
            on the host cores over a bounded sample of the same workload.

`--impl reference` times the reference CPU implementation alone (rank 0 only under torchrun).
Multi-GPU: pictures shard by stream across ranks with no data-path collective ("weak" scaling).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W4K, H4K, BD = 3840, 2160, 10
METRIC = "filtered Mpixel/s (deblock+CDEF+CCSO)"
UNFUSED = os.environ.get("SVTF_CHAIN_UNFUSED") is not None  # A/B: CDEF and CCSO as separate launches
UNIT = "Mpixel/s"


def workload(H, geom, seed, with_mi=True):
    S = H.S
    rec = H.synth_picture(geom, seed)
    soa = H.synth_mode_info(geom, seed + 1, p_split=0.5, p_skip=0.3, p_inter=0.7)
    mi = H.pack_mi(soa) if with_mi else None  # packed words: the product's input (and the oracle port's); the reference reads `soa`
    lf = S.lf_params((24, 24), 12, 12)
    cdef, idx = H.synth_cdef(geom, seed + 2, n_strengths=8, damping=5)
    ccso, flags = H.synth_ccso(geom, seed + 3, quant_idx=(0, 0, 0), support=(0, 1, 2), edge_clf=(0, 0, 0), band_log2=(2, 2, 2), p_on=0.8)
    return dict(rec=rec, soa=soa, mi=mi, lf=lf, cdef=cdef, idx=idx, ccso=ccso, flags=flags)


# ------------------------------------------------------------------------------------------------
# CPU legs
# ------------------------------------------------------------------------------------------------
def bind_to_gpu_cpus(torch, local):
    """Run this rank's host threads (and first-touch its pinned buffers) on the CPUs local to its GPU
    (/sys/bus/pci/devices/<bdf>/local_cpulist): with several ranks per box the end-to-end path is bound by host memory and
    PCIe, and a rank on the wrong socket pays the inter-socket link twice per picture.  Best effort; returns a note."""
    try:
        p = torch.cuda.get_device_properties(local)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/local_cpulist") as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return f"{bdf}: {len(cpus)} local CPUs"
    except Exception as e:  # noqa: BLE001 - purely advisory
        return f"not bound ({type(e).__name__})"
    return "not bound"


def cpu_chain_runner(H, simd=True):
    """Returns (kind, fn(geom, wl) -> None) running deblock->CDEF->CCSO on the CPU with the reference's own code:
    simd=True  oracle/_ref/libsvtref_simd.so — the SSE2 / AVX2 kernels the reference's RTCD selects on this host (its real fast path;
               tests/test_ref_simd.py pins it equal to the C-only build), kind "reference-avx2";
    simd=False oracle/_ref/libsvtref.so — the `_c` kernels, kind "reference (C only, no SIMD)";
    neither library present: the oracle port."""
    if H.ref_available():
        use_simd = simd and H.ref_simd_available()
        ref = H.Ref(simd=use_simd)

        def run(geom, wl):
            a = ref.deblock(geom, wl["rec"], wl["soa"], wl["lf"])
            b = ref.cdef_apply(geom, a, wl["soa"], wl["cdef"], wl["idx"])
            ref.ccso_apply(geom, b, a[0], wl["ccso"], wl["flags"])

        return ("reference-avx2" if use_simd else "reference (C only, no SIMD)"), run
    orc = H.Oracle()

    def run(geom, wl):
        a = orc.deblock(geom, wl["rec"], wl["mi"], wl["lf"])
        b = orc.cdef_apply(geom, a, wl["mi"], wl["cdef"], wl["idx"])
        orc.ccso_apply(geom, b, a[0], wl["ccso"], wl["flags"])

    return "port", run


CPU_SAMPLE_H = 2160  # every host thread filters one whole 3840x2160 10-bit picture per step — the same sample in both arms


def cpu_measure(H, band_h, threads, steps, warmup, simd=True):
    """Each step: `threads` host threads each filter one 3840 x band_h 10-bit picture (ctypes releases the GIL).  Four distinct
    pictures are cycled over the threads (the inputs are read-only; every call filters its own copy)."""
    S = H.S
    geom = S.make_geom(W4K, band_h, BD)
    kind, run = cpu_chain_runner(H, simd)
    nd = min(threads, 4)
    wls = [workload(H, geom, 0xC2 + 16 * t, with_mi=(kind == "port")) for t in range(nd)]

    def step():
        th = [threading.Thread(target=run, args=(geom, wls[t % nd])) for t in range(threads)]
        for t in th:
            t.start()
        for t in th:
            t.join()

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    mpix = threads * W4K * band_h * steps / 1e6
    return kind, mpix / dt, dt / steps, f"{threads} threads x one 3840x{band_h} 10-bit picture per step, {steps} steps"


def cpu_search_measure(H, band_h, threads):
    """CDEF strength search (64 strengths) + pick on the host: `threads` threads, one 3840 x band_h band each —
    the reference's own segment split (EbEncHandle.c:503-506) is by filter-block rectangles like this."""
    S = H.S
    geom = S.make_geom(W4K, band_h, BD)
    f1, f2 = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(f1, f2, base_q_idx=120, lambda_=3000)
    items = []
    for t in range(threads):
        rec = H.synth_picture(geom, 0xC3 + 16 * t)
        soa = H.synth_mode_info(geom, 0xC3 + 16 * t + 1)
        items.append((rec, H.synth_source(geom, rec, 0xC3 + 16 * t + 2), soa, H.pack_mi(soa)))
    if H.ref_available():
        use_simd = H.ref_simd_available()
        ref, kind = H.Ref(simd=use_simd), ("reference-avx2" if use_simd else "reference (C only, no SIMD)")

        def run(it):
            mse, skip, _, _ = ref.cdef_search(geom, it[0], it[1], it[2], cfg)
            ref.cdef_pick(geom, it[2], cfg, mse, skip)
    else:
        orc, kind = H.Oracle(), "port"

        def run(it):
            mse, skip, _, _ = orc.cdef_search(geom, it[0], it[1], it[3], cfg)
            orc.cdef_pick(cfg, mse, skip)
    t0 = time.perf_counter()
    th = [threading.Thread(target=run, args=(it,)) for it in items]
    for t in th:
        t.start()
    for t in th:
        t.join()
    dt = time.perf_counter() - t0
    nfb = threads * ((band_h + 63) // 64) * (W4K // 64)
    return {"value": round(nfb / dt, 1), "unit": "filter blocks/s", "cores": threads, "kind": kind,
            "sample": f"{threads} threads x one 3840x{band_h} 10-bit band, 64 strengths + pick, 1 step"}


def ccso_derive_leg(H, S, geom, wl, ctx, np, with_cpu):
    """CCSO derivation (ccso_search): whole 4K 10-bit picture on the GPU; the reference's C on a small picture."""
    src = H.synth_source(geom, wl["rec"], 0xB200CC, amp=6)
    ctx.begin(geom)
    ctx.upload(wl["rec"])
    ctx.upload_source(src)
    ctx.set_mi(wl["mi"])
    ctx.cdef_frame(wl["cdef"], wl["idx"])
    ctx.sync()
    ctx.ccso_search(4000)  # the first call allocates the derivation's device and pinned buffers
    l0 = ctx.launch_count
    t0 = time.perf_counter()
    out, _ = ctx.ccso_search(4000)
    dt = time.perf_counter() - t0
    res = {"metric": "CCSO derivation Mpixel/s (200 configurations x <=16 training iterations per plane, 3 planes)",
           "value": round(geom.width * geom.height / dt / 1e6, 2), "unit": "Mpixel/s", "ms_per_picture": round(dt * 1e3, 1),
           "gpu_launches": ctx.launch_count - l0, "enabled_planes": [int(out.ccso_enable[p]) for p in range(3)], "cpu_baseline": None}
    if with_cpu and H.ref_available():
        g2 = S.make_geom(480, 272, BD)
        rec = H.synth_picture(g2, 0xCC1)
        org = H.synth_source(g2, rec, 0xCC2, amp=6)
        ref = H.Ref()
        t0 = time.perf_counter()
        for p in range(3):
            ref.ccso_derive_plane(g2, p, rec[0], rec[p], org[p], 4000)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": round(480 * 272 / dt / 1e6, 4), "unit": "Mpixel/s", "cores": 1, "kind": "reference",
                               "sample": "one 480x272 10-bit picture, 3 planes, single thread (derive_ccso_filter is serial)"}
    return res


def lf_search_leg(H, S, geom, wl, ctx, np, with_cpu):
    """Deblocking level search (svt_av1_pick_filter_level, LPF_PICK_FROM_FULL_IMAGE): luma + both chroma planes, each a
    bisection of whole-plane trial deblocks + SSE against the source; the reference's C on a smaller picture."""
    src = H.synth_source(geom, wl["rec"], 0xB200D1, amp=6)
    lf = S.lf_params((24, 24), 16, 16)
    ctx.begin(geom)
    ctx.upload(wl["rec"])
    ctx.upload_source(src)
    ctx.set_mi(wl["mi"])
    ctx.pick_filter_level(lf)  # warm-up
    l0 = ctx.launch_count
    n = 5
    t0 = time.perf_counter()
    for _ in range(n):
        lv = ctx.pick_filter_level(lf)
    dt = (time.perf_counter() - t0) / n
    res = {"metric": "deblock level search Mpixel/s (Y, U, V searched; picture resident)", "value": round(geom.width * geom.height / dt / 1e6, 1),
           "unit": "Mpixel/s", "ms_per_picture": round(dt * 1e3, 3), "gpu_launches_per_picture": (ctx.launch_count - l0) // n,
           "picked_levels": lv, "cpu_baseline": None}
    if with_cpu and H.ref_available():
        g2 = S.make_geom(960, 544, BD)
        rec = H.synth_picture(g2, 0xD11)
        org = H.synth_source(g2, rec, 0xD12, amp=6)
        soa = H.synth_mode_info(g2, 0xD13)
        ref = H.Ref()
        t0 = time.perf_counter()
        ref.lf_search(g2, rec, org, soa, lf)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": round(960 * 544 / dt / 1e6, 2), "unit": "Mpixel/s", "cores": 1, "kind": "reference",
                               "sample": "one 960x544 10-bit picture, single thread (the reference's search is serial per picture)"}
    return res


def superres_leg(H, S, np):
    """Normative super-resolution upscale (svt_av1_superres_upscale_frame) of a picture coded at half width: 1920x2160 -> 3840x2160,
    10-bit, picture resident.  Bandwidth-bound stage: reads b, writes 2b bytes per coded sample (denominator 16)."""
    # the 64 x 8 filter table is the caller's (av1_resize_filter_normative); here the committed fixture (oracle/make_superres_table.py)
    filt = np.load(os.path.join(ROOT, "tests", "golden", "superres_filters.npy"))
    g = S.make_geom(1920, 2160, BD)
    gu = S.make_geom(3840, 2160, BD)
    pic = H.synth_picture(g, 0xB20051)
    with S.FilterContext(gu) as ctx:
        stream = None
        n, ts = 10, []
        for i in range(n + 2):
            ctx.begin(g)
            ctx.upload(pic)
            ctx.sync()
            t0 = time.perf_counter()
            ctx.superres_upscale_frame(3840, 16, filt)
            ctx.sync()
            ts.append(time.perf_counter() - t0)
        dt = sorted(ts[2:])[len(ts[2:]) // 2]
    alg = 1.5 * 1920 * 2160 * 2 * 3  # read b + write 2b per coded sample, b = 2
    return {"metric": "super-resolution upscale 1920x2160 -> 3840x2160 10-bit, ms per picture (call + sync, picture resident)",
            "ms": round(dt * 1e3, 4), "alg_bytes": int(alg), "alg_GBps": round(alg / dt / 1e9, 1),
            "note": "includes the 6 KB table upload and a stream synchronisation per call"}


def lr_search_leg(H, S, geom, wl, ctx, np, with_cpu, more_ctxs=()):
    """Loop-restoration filter search (restoration_seg_search + rest_finish_search), Wiener tools first, then with the self-guided tools: statistics of every unit,
    host solve, refinement trials of all units in lock-step; the reference's C on a smaller picture."""
    src = H.synth_source(geom, wl["rec"], 0xB200F1, amp=6)
    cfg = H.lr_search_cfg(unit_size=(256, 128, 128))
    ctx.begin(geom)
    ctx.upload(src)          # the picture searched: "reconstruction" = source + coding noise
    ctx.upload_source(wl["rec"])
    ctx.restoration_search(cfg)  # warm-up
    l0 = ctx.launch_count
    n = 3
    t0 = time.perf_counter()
    for _ in range(n):
        ftype, units, search = ctx.restoration_search(cfg)
    dt = (time.perf_counter() - t0) / n
    wn_launches = (ctx.launch_count - l0) // n
    t0 = time.perf_counter()
    for _ in range(n):
        ctx.compute_stats(0, 256, 7)
    dts = (time.perf_counter() - t0) / n
    cfg2 = H.lr_search_cfg(unit_size=(256, 128, 128), sg=1)
    ctx.restoration_search(cfg2)
    l1 = ctx.launch_count
    t0 = time.perf_counter()
    for _ in range(n):
        ftype2, _, _ = ctx.restoration_search(cfg2)
    dt2 = (time.perf_counter() - t0) / n
    both = {"metric": "same with the self-guided tools too (16 parameter sets, coefficient refinement, RESTORE_SWITCHABLE)",
            "value": round(geom.width * geom.height / dt2 / 1e6, 1), "unit": "Mpixel/s", "ms_per_picture": round(dt2 * 1e3, 3),
            "gpu_launches_per_picture": (ctx.launch_count - l1) // n, "frame_types": ftype2}
    res = {"metric": "loop-restoration search Mpixel/s (Wiener 7-tap luma + 5-tap chroma, refinement on; picture resident)",
           "value": round(geom.width * geom.height / dt / 1e6, 1), "unit": "Mpixel/s", "ms_per_picture": round(dt * 1e3, 3),
           "gpu_launches_per_picture": wn_launches, "luma_stats_ms_incl_download": round(dts * 1e3, 3),
           "frame_types": ftype, "wiener_and_self_guided": both, "cpu_baseline": None}
    if more_ctxs:  # several pictures in flight: one context + submitting host thread each (the encoder keeps pictures in flight)
        from concurrent.futures import ThreadPoolExecutor

        pool_ctxs = [ctx] + list(more_ctxs)
        for c in pool_ctxs[1:]:
            c.begin(geom)
            c.upload(src)
            c.upload_source(wl["rec"])
        with ThreadPoolExecutor(len(pool_ctxs)) as pool:
            list(pool.map(lambda c: c.restoration_search(cfg), pool_ctxs))
            t0 = time.perf_counter()
            list(pool.map(lambda c: [c.restoration_search(cfg) for _ in range(n)], pool_ctxs))
            dtp = (time.perf_counter() - t0) / (n * len(pool_ctxs))
        res["pictures_in_flight"] = {"n": len(pool_ctxs), "value": round(geom.width * geom.height / dtp / 1e6, 1), "unit": "Mpixel/s",
                                     "ms_per_picture": round(dtp * 1e3, 3)}
    if with_cpu and H.ref_available():
        g2 = S.make_geom(960, 544, BD)
        rec = H.synth_picture(g2, 0xF11)
        org = H.synth_source(g2, rec, 0xF12, amp=6)
        vis = lambda pl: [np.ascontiguousarray(p[: (g2.height + s) >> s, : (g2.width + s) >> s]) for p, s in zip(pl, (0, 1, 1))]
        ref = H.Ref()
        t0 = time.perf_counter()
        ref.lr_search(g2, vis(org), vis(rec), cfg)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": round(960 * 544 / dt / 1e6, 3), "unit": "Mpixel/s", "cores": 1, "kind": "reference",
                               "sample": "one 960x544 10-bit picture, one segment, single thread"}
        g3 = S.make_geom(480, 272, BD)
        rec = H.synth_picture(g3, 0xF13)
        org = H.synth_source(g3, rec, 0xF14, amp=6)
        vis3 = lambda pl: [np.ascontiguousarray(p[: (g3.height + s) >> s, : (g3.width + s) >> s]) for p, s in zip(pl, (0, 1, 1))]
        t0 = time.perf_counter()
        ref.lr_search(g3, vis3(org), vis3(rec), cfg2)
        dt = time.perf_counter() - t0
        both["cpu_baseline"] = {"value": round(480 * 272 / dt / 1e6, 3), "unit": "Mpixel/s", "cores": 1, "kind": "reference",
                                "sample": "one 480x272 10-bit picture, one segment, single thread"}
    return res


def side_info_leg(H, S, geom, ctx, np):
    """Side info to the device: host packer + 4 B per 4x4 over PCIe (svtf_pack_mi_enc + svtf_set_mi) against byte planes
    + packing on the device (svtf_set_mi_soa_enc, 1.5 B per 4x4)."""
    soa = H.synth_mode_info(geom, 0xB200E1)
    keys = ("bsize", "tx_depth", "skip", "ref_frame0", "mode", "segment_id")
    pins = {k: S.PinnedArray(soa[k].shape, np.uint8) for k in keys}
    for k in keys:
        pins[k].array[...] = soa[k]
    psoa = {k: pins[k].array for k in keys}
    pin_mi = S.PinnedArray(soa["bsize"].shape, np.uint32)
    ctx.begin(geom)
    n = 10

    def timed(fn):
        fn()
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        ctx.sync()
        return (time.perf_counter() - t0) / n * 1e3

    def host_path():
        pin_mi.array[...] = H.pack_mi(psoa)
        ctx.set_mi(pin_mi.array)

    t_host = timed(host_path)
    t_up = timed(lambda: ctx.set_mi(pin_mi.array))
    t_soa = timed(lambda: ctx.set_mi_soa(psoa))
    same = bool(np.array_equal(ctx.get_mi(), H.pack_mi(soa)))
    for k in keys:
        pins[k].close()
    pin_mi.close()
    return {"metric": "side info of one 4K picture to the device, ms", "host_pack_plus_upload_ms": round(t_host, 3), "packed_upload_only_ms": round(t_up, 3),
            "byte_planes_plus_device_pack_ms": round(t_soa, 3), "device_words_equal_host_words": same,
            "bytes_per_4x4": {"packed": 4, "byte_planes": 6}}


def lr_leg(H, S, geom, wl, ctx, np, with_cpu):
    """Loop restoration (Wiener / self-guided, 64x64 luma units, mixed types) after deblock + CDEF on the resident picture."""
    lr, units = H.synth_lr(geom, 0xB200F1)
    lf = S.lf_params((24, 24), 16, 16)
    ctx.begin(geom)
    ctx.upload(wl["rec"])
    ctx.set_mi(wl["mi"])
    n = 10

    def run(with_lr):
        ctx.rewind()
        ctx.loop_filter_frame(lf)
        ctx.cdef_frame(wl["cdef"], wl["idx"])
        if with_lr:
            ctx.loop_restoration_filter_frame(lr, units)

    def timed(with_lr):
        run(with_lr)
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(n):
            run(with_lr)
        ctx.sync()
        return (time.perf_counter() - t0) / n

    base, full = timed(False), timed(True)
    dt = max(full - base, 1e-9)
    res = {"metric": "loop restoration Mpixel/s (after deblock + CDEF, picture resident; includes the upload of the unit parameters)",
           "value": round(geom.width * geom.height / dt / 1e6, 1), "unit": "Mpixel/s", "ms_per_picture": round(dt * 1e3, 3),
           "chain_with_lr_ms": round(full * 1e3, 3), "cpu_baseline": None}
    if with_cpu and H.ref_available():
        g2 = S.make_geom(960, 544, BD)
        dbk = H.synth_picture(g2, 0xF11)
        cur = H.synth_source(g2, dbk, 0xF12, amp=5)
        lr2, u2 = H.synth_lr(g2, 0xF13)
        ref = H.Ref()
        t0 = time.perf_counter()
        ref.lr_apply(g2, cur, dbk, lr2, u2)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": round(960 * 544 / dt / 1e6, 2), "unit": "Mpixel/s", "cores": 1, "kind": "reference",
                               "sample": "one 960x544 10-bit picture incl. save_boundary_lines, single thread"}
    return res


def handoff_leg(H, S, geom, wl, ctx, np, with_cpu):
    """Recon hand-off: the filtered picture leaves the device as the padded reference picture
    (pad_ref_and_set_flags: borders of 160 luma samples, 8-bit + 2-bit split) into pinned host memory."""
    ox = oy = 160
    W, Hh = geom.width, geom.height
    ctx.begin(geom)
    ctx.upload(wl["rec"])
    shp = ctx.ref_padded_shapes(W, Hh, ox, oy)
    pins = [[S.PinnedArray(sh, dt) for sh in shp] for dt in (np.uint16, np.uint8, np.uint8)]
    out = tuple([p.array for p in grp] for grp in pins)
    n = 8

    def timed(split):
        o = out if split else (out[0], None, None)
        ctx.download_ref_padded(W, Hh, ox, oy, 0, 0, split, split, out=o)
        t0 = time.perf_counter()
        for _ in range(n):
            ctx.download_ref_padded(W, Hh, ox, oy, 0, 0, split, split, out=o)
        return (time.perf_counter() - t0) / n

    plain, full = timed(False), timed(True)
    samples = sum(a * b for a, b in shp)
    # the kernel alone (padded planes + split stay on the device), CUDA events on the context's stream; the 12 resident
    # pictures of the main leg were written since, so the source surface is not in L2
    import torch

    st = torch.cuda.Stream()
    ctx.set_stream(st.cuda_stream)
    evs = []
    for _ in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        ctx.ref_padded_resident(W, Hh, ox, oy, True, True)
        e1.record(st)
        evs.append((e0, e1))
    st.synchronize()
    ctx.set_stream(None)
    k_ms = statistics.median(a.elapsed_time(b) for a, b in evs[1:])
    src_samples = sum(((geom.mi_cols * 4) >> s) * ((geom.mi_rows * 4) >> s) for s in (0, 1, 1))
    alg = src_samples * 2 + samples * 4  # read every source sample once; write 16-bit + 8-bit + 2-bit planes of the padded picture
    res = {"metric": "recon hand-off ms per 4K 10-bit picture (padded reference picture to pinned host memory, synchronous)",
           "padded_only_ms": round(plain * 1e3, 3), "padded_plus_8bit_2bit_split_ms": round(full * 1e3, 3),
           "d2h_bytes": {"padded_only": samples * 2, "with_split": samples * 4},
           "kernel": {"name": "handoff_kernel", "ms": round(k_ms, 4), "alg_bytes": alg, "alg_GBps": round(alg / (k_ms * 1e-3) / 1e9, 1),
                      "note": "bandwidth-bound stage: read b, write b+1+1 bytes per padded sample; back-to-back launches on one 4K picture (source 25 MB may sit in L2)"},
           "value": round(W * Hh / full / 1e6, 1), "unit": "Mpixel/s", "cpu_baseline": None}
    if with_cpu and H.ref_available():
        ref = H.Ref()
        t0 = time.perf_counter()
        H.handoff_cpu(ref.lib, "ref_", geom, wl["rec"], W, Hh, ox, oy, 0, 0, True, True)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": round(W * Hh / dt / 1e6, 1), "unit": "Mpixel/s", "cores": 1, "kind": "reference",
                               "sample": "one 4K 10-bit picture: copy into the padded buffer + generate_padding16_bit + un_pack2d, single thread"}
    for grp in pins:
        for p in grp:
            p.close()
    return res


def streams_1080p_leg(H, S, torch, dist, dev, local, rank, world, steps):
    """BASELINE configs[4] (and [0]'s picture format): independent 1080p 8-bit streams sharded over the GPUs by stream — 64 streams on
    8 GPUs = 8 per GPU, so every rank runs 8 contexts, one picture of each stream per step, each on its own CUDA stream; no collective.
    Device-resident (the pictures are rewound, not re-uploaded); 8 pictures x 3 surfaces x 3.1 MB fit in L2, so the kernels of this
    leg read from L2 — it measures launch / tail behaviour at the small picture size, not HBM."""
    g = S.make_geom(1920, 1080, 8, 1)
    per_gpu = 8
    wls = [workload(H, g, 0xC5 + 31 * rank + 5 * i) for i in range(2)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(per_gpu)]
    ctxs = []
    for i in range(per_gpu):
        c = S.FilterContext(g, device=local)
        c.begin(g)
        c.upload(wls[i % 2]["rec"])
        c.set_mi(wls[i % 2]["mi"])
        c.sync()
        c.set_stream(streams[i].cuda_stream)
        ctxs.append(c)

    def step():
        for i, c in enumerate(ctxs):
            wl = wls[i % 2]
            c.rewind()
            c.filter_chain(wl["lf"], wl["cdef"], wl["idx"], wl["ccso"], wl["flags"])

    for _ in range(3):
        step()
    torch.cuda.synchronize(dev)
    if dist is not None:
        dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(streams[0])
    for st in streams[1:]:
        st.wait_event(ev0)
    for _ in range(steps):
        step()
    for st in streams[1:]:
        e = torch.cuda.Event()
        e.record(st)
        streams[0].wait_event(e)
    ev1.record(streams[0])
    torch.cuda.synchronize(dev)
    tt = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms = float(tt.item())
    for c in ctxs:
        c.close()
    return {"metric": "filtered Mpixel/s (deblock+CDEF+CCSO), independent 1080p 8-bit streams sharded by stream",
            "value": round(world * per_gpu * 1920 * 1080 * steps / (ms * 1e-3) / 1e6, 1), "unit": "Mpixel/s",
            "streams_per_gpu": per_gpu, "streams_total": world * per_gpu, "pictures_per_s": round(world * per_gpu * steps / (ms * 1e-3), 1),
            "ms_per_step": round(ms / steps, 4), "note": "device-resident, L2-resident working set (24 x 3.1 MB per GPU)"}


def stripes_8k_leg(H, S, torch, dist, dev, local, rank, world, np, steps):
    """BASELINE configs[3]: ONE 8K (7680x4320) 10-bit picture split over the N ranks by superblock-row stripes, strong scaling (the
    picture is fixed as N grows; N = 1 runs the same code on one stripe).  Every rank holds ONLY the rows it owns (svtf_upload_frame_own,
    untimed: the picture is resident as stripes); the timed picture = svtf_stripe_exchange (halo rows from the neighbours' GPUs, NCCL
    send / recv over NVLink) + deblock + CDEF+CCSO on the rank's rows.  Search: per-rank svtf_cdef_search_frame on its rows, then
    svtf_cdef_search_gather_pick (device-to-device gather of the mse / skip rows + the pick on every rank).  Times are CUDA events
    (apply) / wall clock between device synchronisations (search + gather + pick, which ends in a result download), max over ranks."""
    g = S.make_geom(7680, 4320, BD)
    npx = 7680 * 4320
    nv, nh = H.nfb_of(g)
    parts = S.stripes.partition(S.stripes.sb_rows_of(g), world)
    first, count = parts[rank]
    # the same synthetic picture on every rank (each uploads its own rows of it): quarter-size tile repeated 2 x 2 keeps generation short
    gq = S.make_geom(3840, 2160, BD)
    wq = workload(H, gq, 0xB200C4)
    tile2 = lambda a: np.ascontiguousarray(np.tile(a, (2, 2)))
    rec = [tile2(p) for p in wq["rec"]]
    mi = tile2(wq["mi"])
    idx = np.ascontiguousarray(np.tile(wq["idx"].reshape(34, 60), (2, 2)).reshape(-1)[: nv * nh]) if wq["idx"].size == 34 * 60 else None
    cdef, idx2 = H.synth_cdef(g, 0xB200C6, n_strengths=8, damping=5)
    idx = idx2 if idx is None else idx
    ccso, flags = H.synth_ccso(g, 0xB200C7, quant_idx=(0, 0, 0), support=(0, 1, 2), edge_clf=(0, 0, 0), band_log2=(2, 2, 2), p_on=0.8)
    src = [tile2(p) for p in H.synth_source(gq, wq["rec"], 0xB200C5)]
    f1, f2 = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(f1, f2, base_q_idx=120, lambda_=3000)
    # communicator of the library: rank 0 draws the id, torch.distributed carries the 128 bytes
    uid = [S.Comm.unique_id() if rank == 0 else None]
    if dist is not None:
        dist.broadcast_object_list(uid, src=0)
    comm = S.Comm(local, world, rank, uid[0])
    up, down = (rank - 1 if rank > 0 else -1), (rank + 1 if rank + 1 < world else -1)
    NP = 4  # pictures in flight in the pipelined mode
    streams = [torch.cuda.Stream(device=dev) for _ in range(NP)]
    ctxs = []
    for i in range(NP):
        c = S.FilterContext(S.stripes.stripe_geom(g, first, count), device=local)
        c.begin_stripe(g, first, count)
        c.upload_own(rec)
        c.set_mi_own(mi)
        c.upload_source(src)
        c.sync()
        c.set_stream(streams[i].cuda_stream)
        ctxs.append(c)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def one(c):
        c.rewind()
        c.stripe_exchange(comm, up, down)
        c.filter_chain(wq["lf"], cdef, idx, ccso, flags)

    def timed(n_ctx, reps):
        """reps passes over the first n_ctx contexts; returns ms per picture (max over ranks)."""
        for c in ctxs[:n_ctx]:
            one(c)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(streams[0])
        for st in streams[1:n_ctx]:
            st.wait_event(e0)
        for _ in range(reps):
            for c in ctxs[:n_ctx]:
                one(c)
        for st in streams[1:n_ctx]:
            e = torch.cuda.Event()
            e.record(st)
            streams[0].wait_event(e)
        e1.record(streams[0])
        barrier()
        t = torch.tensor([e0.elapsed_time(e1) / (reps * n_ctx)], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    lat_ms = timed(1, max(steps, 10))
    pipe_ms = timed(NP, max(steps // 2, 5))

    def search_once(c):
        c.cdef_seg_search(cfg, want_outputs=False)
        return c.search_gather_pick(comm, cfg, parts, nv * nh)

    ctxs[0].rewind()
    ctxs[0].loop_filter_frame(wq["lf"])
    search_once(ctxs[0])
    barrier()
    t0 = time.perf_counter()
    reps = max(steps // 2, 5)
    for _ in range(reps):
        pick = search_once(ctxs[0])
    barrier()
    s_ms = (time.perf_counter() - t0) / reps * 1e3
    t = torch.tensor([s_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    s_ms = float(t.item())
    for c in ctxs:
        c.close()
    comm.close()
    return {"metric": "one 8K (7680x4320) 10-bit picture split by superblock-row stripes over the ranks (strong scaling), stripes resident",
            "n_gpus": world, "pictures_per_step": 1, "apply_ms_per_picture": round(lat_ms, 4), "apply_Mpixel_per_s": round(npx / lat_ms / 1e3, 1),
            "apply_pipelined": {"pictures_in_flight": NP, "ms_per_picture": round(pipe_ms, 4), "Mpixel_per_s": round(npx / pipe_ms / 1e3, 1)},
            "search_gather_pick_ms_per_picture": round(s_ms, 3), "search_fb_per_s": round(nv * nh / s_ms * 1e3, 1),
            "picked_cdef_bits": int(pick[0].cdef_bits),
            "includes": "apply: NCCL halo exchange (8 luma / 4 chroma / 2 mode-info rows per interior boundary, device to device) + deblock + CDEF+CCSO; "
                        "search: search kernel on the rank's rows + NCCL table gather (device to device) + pick + result download"}


def adapter_e2e_leg(H, S, geom, wls, nsrc, B, steps, barrier, device, np):
    """The headline end-to-end number: the chain through the reference-facing PLUGIN call, i.e. the reference's own entry points
        svt_av1_loop_filter_frame(recon, pcs, 0, 3) -> svt_av1_cdef_frame(scs, pcs) -> ccso_frame(recon, pcs, pd, ext_rec_y)
    (adapters/svtf_enc_adapters.c) on PictureControlSet structures with an ordinary malloc'ed recon picture and a real pointer grid
    (mi_grid_base -> ModeInfo), as the pipeline's DLF / CDEF stages would call them: per picture one gather of the mode info, one upload,
    three launches, one download into the same picture buffers.  B pictures per pass, 4 host threads (the reference runs its filter
    stages on several pictures at once the same way).  Between passes the pictures are reloaded with the unfiltered reconstruction
    (what EncDec leaves there) OUTSIDE the timed region.  Returns None where the adapter test library is absent."""
    import ctypes as C
    from concurrent.futures import ThreadPoolExecutor

    path = os.path.join(ROOT, "adapters", "_build", "libsvtadapt_test.so")
    if not os.path.exists(path):
        return None
    lib = C.CDLL(path)
    lib.adapt_chain_open.restype = C.c_void_p
    lib.adapt_chain_open.argtypes = [C.c_void_p] * 7
    lib.adapt_chain_run.argtypes = [C.c_void_p]
    lib.adapt_chain_close.argtypes = [C.c_void_p]
    lib.adapt_chain_load.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    lib.adapt_chain_planes.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.svtf_adapter_set_device(device)
    NT = int(os.environ.get("SVTF_BENCH_ADAPTER_THREADS", "6"))   # pictures in flight = host threads calling the entry points
    GT = int(os.environ.get("SVTF_BENCH_GATHER_THREADS", "4"))    # threads gathering one picture's mode info
    lib.adapt_set_modes(1, 1, GT)  # resident chain, page-locked picture buffers
    keep, handles = [], []
    for i in range(B):
        wl = wls[i % nsrc]
        soa, k = H.Ref._soa(wl["soa"])
        fl = (C.c_void_p * 3)(*[f.ctypes.data for f in wl["flags"]])
        idx = np.ascontiguousarray(wl["idx"], dtype=np.int8)
        keep += [soa, k, fl, idx]
        h = lib.adapt_chain_open(C.addressof(geom), C.addressof(soa), C.addressof(wl["lf"]), C.addressof(wl["cdef"]), idx.ctypes.data, C.addressof(wl["ccso"]), fl)
        assert h, "adapt_chain_open failed"
        handles.append(h)

    def load_all():
        for i, h in enumerate(handles):
            pl = wls[i % nsrc]["rec"]
            lib.adapt_chain_load(h, *H._planes(pl))

    pool = ThreadPoolExecutor(NT)

    def run_all():
        def worker(t):
            for h in handles[t::NT]:
                lib.adapt_chain_run(h)
        list(pool.map(worker, range(NT)))

    load_all()
    run_all()  # warm-up: contexts, page locks, staging
    # check one picture against the device-resident library result
    outp, outs = (C.c_void_p * 3)(), (C.c_int * 3)()
    lib.adapt_chain_planes(handles[0], outp, outs)
    wl = wls[0]
    with S.FilterContext(geom, device=device) as c:
        c.begin(geom)
        c.upload(wl["rec"])
        c.set_mi(wl["mi"])
        c.filter_chain(wl["lf"], wl["cdef"], wl["idx"], wl["ccso"], wl["flags"])
        exp = c.download()
    ok = True
    for p in range(3):
        hh, ww = exp[p].shape
        buf = np.ctypeslib.as_array(C.cast(outp[p], C.POINTER(C.c_uint16 if geom.sample_bytes == 2 else C.c_uint8)), shape=(hh, outs[p]))
        ok = ok and bool(np.array_equal(buf[:, :ww], exp[p]))
    secs, passes = 0.0, max(3, steps)
    for _ in range(passes):
        load_all()
        barrier()
        t0 = time.perf_counter()
        run_all()
        secs += time.perf_counter() - t0
    pool.shutdown()
    for h in handles:
        lib.adapt_chain_close(h)
    return {"seconds": secs, "pictures": B * passes, "host_threads": NT, "matches_device_path": ok, "ms_per_picture": round(secs / (B * passes) * 1e3, 4)}


def app_leg(frames):
    """BASELINE configs[0] for context: the reference's own SvtAv1EncApp --preset 8 on a 1080p 8-bit clip, unmodified and with the filter
    stages replaced by the adapters over libsvtfilt_b200.so (integration/build.sh builds both; C-only builds, the image has no nasm), then
    both SvtAv1DecApp builds on the resulting stream.  Reports frames per second of the four runs and whether reconstruction and bitstream
    are byte-identical — the applications are dominated by mode decision on the host, so the fps are context, not a result."""
    import hashlib
    import subprocess
    import tempfile

    bins = os.path.join(ROOT, "integration", "_bin")
    need = ["SvtAv1EncApp_ref", "SvtAv1EncApp_b200", "SvtAv1DecApp_ref", "SvtAv1DecApp_b200"]
    if not all(os.path.exists(os.path.join(bins, b)) for b in need):
        return None
    sys.path.insert(0, os.path.join(ROOT, "integration"))
    import synth_yuv

    def md5(path):
        h = hashlib.md5()
        with open(path, "rb") as f:
            for blk in iter(lambda: f.read(1 << 22), b""):
                h.update(blk)
        return h.hexdigest()

    w, h = 1920, 1080
    out = {"workload": f"SvtAv1EncApp --preset 8, {w}x{h} 8-bit, {frames} synthetic frames; SvtAv1DecApp on that stream", "build": "C only (no nasm in the image)"}
    with tempfile.TemporaryDirectory() as td:
        clip = os.path.join(td, "c.yuv")
        synth_yuv.write_clip(clip, w, h, frames, 8)
        for tag in ("ref", "b200"):
            cmd = [os.path.join(bins, "SvtAv1EncApp_" + tag), "-i", clip, "-w", str(w), "-h", str(h), "--fps", "30", "-n", str(frames), "--preset", "8",
                   "--input-depth", "8", "-b", os.path.join(td, tag + ".ivf"), "-o", os.path.join(td, tag + ".yuv")]
            t0 = time.perf_counter()
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
            out["enc_fps_" + tag] = round(frames / (time.perf_counter() - t0), 2) if r.returncode == 0 else None
        ok = all(out["enc_fps_" + t] for t in ("ref", "b200"))
        out["recon_identical"] = ok and md5(os.path.join(td, "ref.yuv")) == md5(os.path.join(td, "b200.yuv"))
        out["bitstream_identical"] = ok and md5(os.path.join(td, "ref.ivf")) == md5(os.path.join(td, "b200.ivf"))
        if ok:
            for tag in ("ref", "b200"):
                t0 = time.perf_counter()
                r = subprocess.run([os.path.join(bins, "SvtAv1DecApp_" + tag), "-i", os.path.join(td, "ref.ivf"), "-o", os.path.join(td, "dec_" + tag + ".yuv")],
                                   capture_output=True, text=True, timeout=900)
                out["dec_fps_" + tag] = round(frames / (time.perf_counter() - t0), 2) if r.returncode == 0 else None
            out["decoded_identical"] = bool(out["dec_fps_ref"] and out["dec_fps_b200"]) and md5(os.path.join(td, "dec_ref.yuv")) == md5(os.path.join(td, "dec_b200.yuv"))
    return out


def pcie_probe_leg(torch, dist, dev, world):
    """What bounds the end-to-end numbers as ranks are added: every rank copies 25 MB blocks (one 4K 10-bit picture) between pinned host
    memory and its GPU — H2D alone, D2H alone, both at once — ALL RANKS AT THE SAME TIME; per-rank GB/s from the slowest rank's time and
    the box total.  If the total stops growing with the ranks, the host side of the box (memory / root complex), not the GPUs, is the
    limit of e2e."""
    mb, n = 25, 24
    hin = [torch.empty(mb << 20, dtype=torch.uint8).pin_memory() for _ in range(4)]
    hout = [torch.empty(mb << 20, dtype=torch.uint8).pin_memory() for _ in range(4)]
    d = [torch.empty(mb << 20, dtype=torch.uint8, device=dev) for _ in range(4)]
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(h2d, d2h):
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s1)
        f0.record(s2)
        for i in range(n):
            if h2d:
                with torch.cuda.stream(s1):
                    d[i % 2].copy_(hin[i % 4], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    hout[i % 4].copy_(d[2 + i % 2], non_blocking=True)
        e1.record(s1)
        f1.record(s2)
        torch.cuda.synchronize(dev)
        t = torch.tensor([max(e0.elapsed_time(e1), f0.elapsed_time(f1))], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return n * (mb << 20) / (float(t.item()) * 1e-3) / 1e9

    run(True, True)
    h, dd, dup = run(True, False), run(False, True), run(True, True)
    return {"block_MB": mb, "ranks_at_once": world, "per_rank_GBps": {"h2d": round(h, 1), "d2h": round(dd, 1), "duplex_each_way": round(dup, 1)},
            "box_total_GBps": {"h2d": round(h * world, 1), "d2h": round(dd * world, 1), "duplex_each_way": round(dup * world, 1)}}


def gpu_search_leg(H, S, geom, wl, ctxs, steps, np):
    """Metric (ii): filter blocks per second through cdef_seg_search (64 strengths, all planes) + finish_cdef_search.
    Several pictures are in flight, one context + stream + submitting host thread each — the reference's CDEF stage runs
    one thread per picture the same way — so the short dependent launches of one picture's pick overlap another's search."""
    from concurrent.futures import ThreadPoolExecutor

    f1, f2 = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(f1, f2, base_q_idx=120, lambda_=3000)
    src = H.synth_source(geom, wl["rec"], 0xB200C3)
    nfb = ((geom.mi_rows + 15) // 16) * ((geom.mi_cols + 15) // 16)
    pin_rec, pin_src = S.PinnedPicture(geom), S.PinnedPicture(geom)
    for p in range(3):
        pin_rec.planes[p][...] = wl["rec"][p]
        pin_src.planes[p][...] = src[p]
    pin_mi = S.PinnedArray(wl["mi"].shape, np.uint32)
    pin_mi.array[...] = wl["mi"]

    def load(ctx):
        ctx.begin(geom)
        ctx.upload(pin_rec.planes)
        ctx.upload_source(pin_src.planes)
        ctx.set_mi(pin_mi.array)

    def resident(ctx):
        ctx.cdef_seg_search(cfg, want_outputs=False)
        return ctx.finish_cdef_search(cfg)  # ends with the D2H of the picked strengths + per-fb indices (sync)

    def e2e(ctx):
        load(ctx)
        return resident(ctx)

    for c in ctxs:
        load(c)
        c.sync()
    out = {}
    with ThreadPoolExecutor(len(ctxs)) as pool:
        for name, fn in (("value", resident), ("e2e", e2e)):
            def worker(c, n):
                for _ in range(n):
                    r = fn(c)
                return r
            list(pool.map(lambda c: worker(c, 2), ctxs))
            t0 = time.perf_counter()
            picks = list(pool.map(lambda c: worker(c, steps), ctxs))
            out[name] = round(nfb * steps * len(ctxs) / (time.perf_counter() - t0), 1)
    t0 = time.perf_counter()
    for _ in range(steps):
        resident(ctxs[0])
    single = round(nfb * steps / (time.perf_counter() - t0), 1)
    # the search kernel alone (device time between two syncs; launch included), for the issue roofline below
    ctxs[0].sync()
    t0 = time.perf_counter()
    for _ in range(steps):
        ctxs[0].cdef_seg_search(cfg, want_outputs=False)
    ctxs[0].sync()
    kernel_ms = (time.perf_counter() - t0) / steps * 1e3
    pin_rec.close()
    pin_src.close()
    pin_mi.close()
    return {"metric": "CDEF-search filter blocks/s (64 strengths, 3 planes, + finish_cdef_search)", "value": out["value"], "e2e": out["e2e"],
            "unit": "filter blocks/s", "pictures_in_flight": len(ctxs), "one_picture_at_a_time": single, "search_kernel_ms": round(kernel_ms, 4),
            "workload": "4K 10-bit, cdef_level 1 list (16 primary + 48 secondary), subsampling 1",
            "picked_cdef_bits": int(picks[0][0].cdef_bits)}


def search_roofline(search, geom, sms, clocks):
    """SURVEY 8(d) for the search: it is compute-bound (93 tap evaluations per compulsory byte), so the roofline is instruction issue.
    Live: the search kernel's time (this run) -> warp instructions per second against SMs x 4 schedulers x the SM clock sampled during the
    run, and the reference-equivalent tap evaluations per second (a separable kernel is credited with what cdef_seg_search would do:
    samples x strengths x 12).  From the committed ncu capture of the same kernel (profiles/search_roofline.json): the instruction count
    per launch and the issue-slot / ALU-pipe / FMA-pipe utilisation."""
    path = os.path.join(ROOT, "profiles", "search_roofline.json")
    ncu = json.load(open(path)) if os.path.exists(path) else {}
    nfb = ((geom.mi_rows + 15) // 16) * ((geom.mi_cols + 15) // 16)
    samples = geom.mi_rows * 4 * geom.mi_cols * 4 * 3 // 2
    tap_evals = samples * 64 * 12
    t = search["search_kernel_ms"] * 1e-3
    mhz = (clocks or {}).get("sm_mhz") or ncu.get("sm_mhz_assumed", 1965)
    peak = sms * 4 * mhz * 1e6
    wi = ncu.get("warp_instructions_per_launch")
    return {"bound": "issue", "kernel": "cdef_search3_kernel", "search_kernel_ms": search["search_kernel_ms"],
            "warp_instructions_per_launch": wi, "achieved": round(wi / t / 1e9, 1) if wi else None, "peak": round(peak / 1e9, 1),
            "unit": "G warp-instructions/s", "frac": round(wi / t / peak, 4) if wi else None,
            "ncu": {k: ncu.get(k) for k in ("issue_active_pct", "alu_pipe_pct", "fma_pipe_pct", "lsu_pipe_pct", "source")},
            "reference_equivalent_tap_evals_per_picture": tap_evals,
            "tap_evals_per_s_kernel": round(tap_evals / t / 1e9, 1), "tap_evals_per_s_with_pick_pipelined": round(search["value"] / nfb * tap_evals / 1e9, 1),
            "tap_unit": "G tap evaluations/s (samples x 64 strengths x 12 taps, what cdef_seg_search evaluates)"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # the reference's CPU code alone: this process neither imports the product's Python package nor maps libsvtfilt_b200.so
    os.environ["SVTF_HARNESS_NO_PRODUCT"] = "1"
    import harness as H

    threads = os.cpu_count() or 1
    kind, val, per_step, sample = cpu_measure(H, CPU_SAMPLE_H, threads, args.steps, args.warmup, simd=True)
    c_only = None
    if kind == "reference-avx2":  # the C-only build (the parity checker) beside it, one step
        k2, v2, _, s2 = cpu_measure(H, CPU_SAMPLE_H, threads, 1, 0, simd=False)
        c_only = {"value": round(v2, 3), "unit": UNIT, "cores": threads, "kind": k2, "sample": s2}
    with open("/proc/self/maps") as f:
        product_loaded = "libsvtfilt_b200" in f.read() or "svtfilt_b200" in sys.modules
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(per_step * 1e3, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u16", "data": "synthetic",
        "config": {"workload": "decoder post-recon filter chain, 4K (3840x2160) 10-bit 4:2:0, deblock+CDEF+CCSO", "sample": sample,
                   "product_library_loaded": bool(product_loaded)},
        "cpu_baseline": {"value": round(val, 3), "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "cpu_baseline_c_only": c_only,
        "e2e": {"value": round(val, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_line(line)
    return 0


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,utilization.gpu,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        busy = []
        for ln in self.f.read().splitlines():
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                c, m, u = float(parts[0]), float(parts[1]), float(parts[2])
            except ValueError:
                continue
            sm.append(c)
            mx.append(m)
            if u > 0:
                busy.append(c)
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        self.f.close()
        os.unlink(self.f.name)
        src = busy or sm
        if src:
            out["sm_mhz"] = statistics.median(src)
            out["sm_max_mhz"] = max(mx)
        out["reasons"] = sorted(reasons)
        out["samples"] = len(sm)
        return out


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
_STDOUT_FD = None


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries below this script write there too (NCCL prints its version line from C when a
    communicator is created), so file descriptor 1 is pointed at stderr for the whole run and the line goes out through a saved copy."""
    global _STDOUT_FD
    if _STDOUT_FD is None:
        sys.stdout.flush()
        _STDOUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _STDOUT_FD is None:
        os.write(1, data)
    else:
        os.write(_STDOUT_FD, data)


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=12, help="distinct 4K pictures resident per GPU (12 x 24.9 MB x 3 surfaces > L2)")
    ap.add_argument("--streams", type=int, default=4, help="CUDA streams the resident pictures are spread over (independent pictures)")
    ap.add_argument("--e2e-inflight", type=int, default=4)
    ap.add_argument("--search-inflight", type=int, default=8, help="pictures in flight in the CDEF-search leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-stripes", action="store_true", help="skip the single-8K-picture stripe leg (configs[3])")
    ap.add_argument("--passes", type=int, default=24, help="passes of the chain over the resident batch per step (one step >= 25 ms of GPU work)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference_arm(args)

    import numpy as np
    import torch
    import harness as H

    S = H.S
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # (a single rank keeps every host core: its cpu_baseline leg times the reference on all of them)
    numa = bind_to_gpu_cpus(torch, local) if world > 1 else "all host CPUs (single rank)"

    geom = S.make_geom(W4K, H4K, BD)
    npx = W4K * H4K
    B = args.batch
    # a few distinct synthetic pictures, cycled over the batch with different seeds per rank
    nsrc = min(B, 3)
    wls = [workload(H, geom, 0xB200C2 + 97 * rank + 7 * i) for i in range(nsrc)]

    # Independent pictures are spread over NS streams (one context per picture, several streams per GPU), so
    # the tail wave of one picture's kernel overlaps the next picture's kernels.  Stream 0 is the timing stream: the
    # others fork from ev0 and join before ev1, so ev0..ev1 brackets every kernel of the K steps on the device.
    NS = max(1, min(args.streams, B))
    streams = [torch.cuda.Stream(device=dev) for _ in range(NS)]
    stream = streams[0]
    ctxs = []
    for i in range(B):
        c = S.FilterContext(geom, device=local)
        c.begin(geom)
        wl = wls[i % nsrc]
        c.upload(wl["rec"])
        c.set_mi(wl["mi"])
        c.sync()
        c.set_stream(streams[i % NS].cuda_stream)
        ctxs.append(c)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    P = max(1, args.passes)  # one step = P passes over the resident batch: >= 25 ms of GPU work, a timed region of >= 0.5 s

    def step(rebuild_planes=False):
        for _ in range(P):
            for i, c in enumerate(ctxs):
                wl = wls[i % nsrc]
                c.rewind()
                if rebuild_planes:  # the second timed loop below: the edge-code planes derived again for every picture
                    c.deblock_prepare()
                c.filter_chain(wl["lf"], wl["cdef"], wl["idx"], wl["ccso"], wl["flags"])

    def stage_pass(stage_events):
        """Every kernel alone on the device (all pictures on stream 0's order), CUDA events around each launch."""
        for i, c in enumerate(ctxs):
            wl = wls[i % nsrc]
            c.set_stream(stream.cuda_stream)
            c.rewind()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(4 if UNFUSED else 3)]
            pe = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            pe[0].record(stream)
            c.deblock_prepare()  # what svtf_set_mi* launches behind its copy (outside `value`, inside e2e): timed here for the record
            pe[1].record(stream)
            prep_events.append(pe)
            e[0].record(stream)
            c.loop_filter_frame(wl["lf"])
            e[1].record(stream)
            if UNFUSED:
                c.cdef_frame(wl["cdef"], wl["idx"])
                e[2].record(stream)
                c.ccso_frame(wl["ccso"], wl["flags"])
                e[3].record(stream)
            else:  # CDEF + CCSO: one kernel (CCSO in the epilogue)
                c.filter_chain(None, wl["cdef"], wl["idx"], wl["ccso"], wl["flags"])
                e[2].record(stream)
            stage_events.append(e)
            c.set_stream(streams[i % NS].cuda_stream)

    sampler = ClockSampler(local) if rank == 0 else None
    for _ in range(args.warmup):
        step()
    barrier()
    # per-kernel durations for the roofline: one serialised pass over the batch (12 pictures > L2), each launch timed alone
    stage_events, prep_events = [], []
    stage_pass(stage_events)
    barrier()
    l0 = sum(c.launch_count for c in ctxs)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for st in streams[1:]:
        st.wait_event(ev0)
    for _ in range(args.steps):
        step()
    for st in streams[1:]:
        e = torch.cuda.Event()
        e.record(st)
        stream.wait_event(e)
    ev1.record(stream)
    barrier()
    launches = sum(c.launch_count for c in ctxs) - l0
    ms = ev0.elapsed_time(ev1)
    tt = torch.tensor([ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_max = float(tt.item())
    value = world * B * P * npx * args.steps / (ms_max * 1e-3) / 1e6
    # The same loop with the mode-info-derived edge-code planes rebuilt for every picture inside the timed region (svtf_set_mi* builds
    # them once per upload, outside `value`): reported beside `value` so that both readings of "inputs resident" are measured.
    steps2 = max(2, args.steps // 4)
    barrier()
    ev0.record(stream)
    for st in streams[1:]:
        st.wait_event(ev0)
    for _ in range(steps2):
        step(rebuild_planes=True)
    for st in streams[1:]:
        e = torch.cuda.Event()
        e.record(st)
        stream.wait_event(e)
    ev1.record(stream)
    barrier()
    t2 = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    value_rebuild = world * B * P * npx * steps2 / (float(t2.item()) * 1e-3) / 1e6

    b = geom.sample_bytes
    # SURVEY 8(d) per-stage algorithmic bytes per luma pixel: deblock 3b+0.5, CDEF 3b, CCSO 4b.  The fused CDEF+CCSO launch is scored
    # against the sum of the stages it replaces (3b + 4b = 7b), as 8(d) says; chain_frac is the whole chain against the 8(d) sum.
    survey_bytes = ((3 * b + 0.5) + 3 * b + 4 * b) * npx
    if UNFUSED:
        names = ["deblock_kernel", "cdef5_kernel", "ccso_kernel"]
        alg_bytes = [(3 * b + 0.5) * npx, 3 * b * npx, 4 * b * npx]
    else:
        names = ["deblock_kernel", "cdef5_kernel+ccso"]
        alg_bytes = [(3 * b + 0.5) * npx, (3 * b + 4 * b) * npx]  # a fused launch is scored against the sum of the stages it replaces (8(d))
    nst = len(names)
    stage_ms = [statistics.mean(e[k].elapsed_time(e[k + 1]) for e in stage_events) for k in range(nst)]
    dom = max(range(nst), key=lambda k: stage_ms[k])
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    achieved = alg_bytes[dom] / (stage_ms[dom] * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath)).get(names[dom])
    roofline = {"bound": "hbm", "kernel": names[dom], "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                "stages": {names[k]: {"ms": round(stage_ms[k], 4), "alg_GBps": round(alg_bytes[k] / (stage_ms[k] * 1e-3) / 1e9, 1)} for k in range(nst)},
                "chain_frac": round(survey_bytes / (sum(stage_ms) * 1e-3) / 1e9 / peak, 4),
                "mode_info_prepare": {"kernel": "dbk_edge_plane_kernel", "ms": round(statistics.mean(a.elapsed_time(b) for a, b in prep_events), 4),
                                      "value_with_planes_rebuilt_per_picture": round(value_rebuild, 1),
                                      "note": "once per mode-info upload (launched by svtf_set_mi* behind the copy): part of e2e, not of value, "
                                              "whose inputs - picture and mode info as uploaded - are resident when the timed region starts"}}

    # ---- e2e: host buffers through the C ABI, several pictures in flight ----
    for c in ctxs:
        c.set_stream(None)
    nin = min(args.e2e_inflight, B)
    pins_in = [S.PinnedPicture(geom) for _ in range(nin)]
    pins_out = [S.PinnedPicture(geom) for _ in range(nin)]
    for k in range(nin):
        for p in range(3):
            pins_in[k].planes[p][...] = wls[k % nsrc]["rec"][p]
    # mode-info grids in pinned memory as well (svtf_host_alloc), so svtf_set_mi is an asynchronous DMA
    pins_mi = [S.PinnedArray(wls[k % nsrc]["mi"].shape, np.uint32) for k in range(nin)]
    for k in range(nin):
        pins_mi[k].array[...] = wls[k % nsrc]["mi"]
    L = S._lib.load()
    import ctypes as C

    def e2e_submit(k):
        c, wl = ctxs[k], wls[k % nsrc]
        c.begin(geom)
        c.upload(pins_in[k].planes)
        c.set_mi(pins_mi[k].array)
        c.filter_chain(wl["lf"], wl["cdef"], wl["idx"], wl["ccso"], wl["flags"])

    def e2e_fetch(k):
        # asynchronous download into pinned memory; completion is awaited before the slot is reused
        c = ctxs[k]
        c._chk(L.svtf_download_frame_async(c._h, *c._planes_args(pins_out[k].planes)))

    EP = 3  # passes over the batch per e2e step (>= 25 ms per step)

    def e2e_step():
        for _ in range(EP):
            e2e_pass()

    def e2e_pass():
        # Software-pipelined submission: picture j's download is queued AFTER picture j+1's upload.  The copy queues
        # issue in submission order across streams, so a download queued first (it has to wait for its kernels) would
        # hold back the next upload for the length of the kernels (tools/e2e_timeline.py shows the lock-step).
        for j in range(B):
            k = j % nin
            if j >= nin:
                ctxs[k].sync()
            e2e_submit(k)
            if j >= 1:
                e2e_fetch((j - 1) % nin)
        e2e_fetch((B - 1) % nin)
        for k in range(nin):
            ctxs[k].sync()

    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    tt = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    e2e_tier1_val = world * B * EP * npx * args.steps / float(tt.item()) / 1e6
    # ---- e2e through the PLUGIN call: the reference's own entry points (Tier-2 adapters) on PictureControlSet structures ----
    # B * EP pictures per pass, like the Tier-1 leg: with only 12 the host threads all upload, then all download (every pass starts cold)
    os.environ.setdefault("SVTF_BENCH_ADAPTER_THREADS", "8" if world == 1 else "6")
    adapter = adapter_e2e_leg(H, S, geom, wls, nsrc, B * EP, max(3, args.steps // 2), barrier, local, np)
    e2e_val = e2e_tier1_val
    if adapter is not None:
        ta = torch.tensor([adapter["seconds"]], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(ta, op=dist.ReduceOp.MAX)
        e2e_val = world * adapter["pictures"] * npx / float(ta.item()) / 1e6
    frame_bytes = int(npx * 1.5 * b)
    mi_bytes = int(wls[0]["mi"].nbytes)
    # check the last e2e output against the device-resident result (same picture, same parameters)
    ctxs[0].begin(geom)
    ctxs[0].upload(wls[0]["rec"])
    ctxs[0].set_mi(wls[0]["mi"])
    ctxs[0].filter_chain(wls[0]["lf"], wls[0]["cdef"], wls[0]["idx"], wls[0]["ccso"], wls[0]["flags"])
    chk = ctxs[0].download()
    e2e_ok = all(np.array_equal(chk[p], pins_out[0].planes[p]) for p in range(3))
    clocks = sampler.stop() if sampler else None

    pcie = pcie_probe_leg(torch, dist, dev, world)
    app = app_leg(int(os.environ.get("SVTF_BENCH_APP_FRAMES", "20")))  # configs[0] has 60 frames: SVTF_BENCH_APP_FRAMES=60 (minutes on C-only builds) if rank == 0 and world == 1 and not args.no_cpu else None
    c5 = streams_1080p_leg(H, S, torch, dist, dev, local, rank, world, max(5, args.steps))
    s8k = None if args.no_stripes else stripes_8k_leg(H, S, torch, dist, dev, local, rank, world, np, args.steps)
    search = gpu_search_leg(H, S, geom, wls[0], ctxs[:args.search_inflight], max(3, args.steps // 2), np) if rank == 0 else None
    if search:
        search["roofline"] = search_roofline(search, geom, torch.cuda.get_device_properties(dev).multi_processor_count, clocks)
    derive = ccso_derive_leg(H, S, geom, wls[0], ctxs[0], np, world == 1 and not args.no_cpu) if rank == 0 else None
    lrl = lr_leg(H, S, geom, wls[0], ctxs[0], np, world == 1 and not args.no_cpu) if rank == 0 else None
    side = side_info_leg(H, S, geom, ctxs[0], np) if rank == 0 else None
    hand = handoff_leg(H, S, geom, wls[0], ctxs[0], np, world == 1 and not args.no_cpu) if rank == 0 else None
    lfs = lf_search_leg(H, S, geom, wls[0], ctxs[0], np, world == 1 and not args.no_cpu) if rank == 0 else None
    sres = superres_leg(H, S, np) if rank == 0 else None
    lrs = lr_search_leg(H, S, geom, wls[0], ctxs[0], np, world == 1 and not args.no_cpu, ctxs[1:4]) if rank == 0 else None
    cpu = search_cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        kind, val, per_step, sample = cpu_measure(H, CPU_SAMPLE_H, threads, 3, 1, simd=True)
        cpu = {"value": round(val, 3), "unit": UNIT, "cores": threads, "kind": kind, "sample": sample}
        if kind == "reference-avx2":
            k2, v2, _, s2 = cpu_measure(H, CPU_SAMPLE_H, threads, 1, 0, simd=False)
            cpu["c_only"] = {"value": round(v2, 3), "unit": UNIT, "kind": k2, "sample": s2}
        search_cpu = cpu_search_measure(H, 64, threads)

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms_max / args.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u16", "data": "synthetic",
            "config": {"workload": "decoder post-recon filter chain, 4K (3840x2160) 10-bit 4:2:0, deblock+CDEF+CCSO",
                       "pictures_per_gpu_per_step": B * P, "passes_per_step": P, "l2": f"inputs larger than L2: {B} pictures x 3 surfaces x {frame_bytes / 1e6:.1f} MB resident",
                       "mode_info": "resident as svtf_set_mi leaves it: packed words + deblocking edge-code planes (roofline.mode_info_prepare)",
                       "streams_per_gpu": NS, "host_affinity": numa,
                       "sharding": f"{world} rank(s), pictures sharded by stream, no collective"},
            "e2e": ({"value": round(e2e_val, 1), "unit": UNIT, "h2d_bytes_per_step": (frame_bytes + 6 * mi_bytes // 4) * B * EP, "d2h_bytes_per_step": frame_bytes * B * EP,
                     "pictures_per_step": B * EP, "path": "Tier-2 adapters: svt_av1_loop_filter_frame -> svt_av1_cdef_frame -> ccso_frame on PictureControlSet structures "
                     "(malloc'ed recon picture page-locked in place once, mode info gathered from the mi_grid_base pointer grid per picture)",
                     "host_threads": adapter["host_threads"], "ms_per_picture": adapter["ms_per_picture"], "matches_device_path": adapter["matches_device_path"]}
                    if adapter is not None else
                    {"value": round(e2e_val, 1), "unit": UNIT, "h2d_bytes_per_step": (frame_bytes + mi_bytes) * B * EP, "d2h_bytes_per_step": frame_bytes * B * EP,
                     "pictures_per_step": B * EP, "path": "Tier-1 C ABI (adapter test library absent)", "matches_device_path": bool(e2e_ok)}),
            "e2e_tier1": {"value": round(e2e_tier1_val, 1), "unit": UNIT, "h2d_bytes_per_step": (frame_bytes + mi_bytes) * B * EP, "d2h_bytes_per_step": frame_bytes * B * EP,
                          "pictures_per_step": B * EP, "inflight": nin, "matches_device_path": bool(e2e_ok),
                          "path": "Tier-1 C ABI from pinned host buffers, mode-info words packed outside the timed region"},
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
            "search": dict(search, cpu_baseline=search_cpu) if search else None, "ccso_derive": derive, "lf_search": lfs, "lr_search": lrs, "superres": sres, "side_info": side, "loop_restoration": lrl, "handoff": hand, "streams_1080p": c5, "stripes_8k": s8k, "pcie": pcie, "app": app,
        }
        emit_line(line)
    for p in pins_in + pins_out + pins_mi:
        p.close()
    for c in ctxs:
        c.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
