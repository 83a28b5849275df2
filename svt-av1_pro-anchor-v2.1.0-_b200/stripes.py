"""Single-picture split across GPUs by superblock-row stripes.

One process per GPU.  Rank g owns superblock rows partition(sb_rows, world)[g]; its FilterContext is put in stripe
mode (svtf_stripe_begin) and is then driven with the same whole-picture arguments as every other rank.  The apply
chain needs no exchange between ranks (the 8 boundary rows are part of each rank's upload).  The CDEF strength search
has one real exchange: every rank holds its filter-block rows of mse/skip, finish_cdef_search needs all of them ->
one all_gather (NCCL on GPUs, gloo in the CPU tests), after which every rank runs the same deterministic pick."""
import numpy as np

from . import filters as F


def sb_rows_of(geom):
    return (geom.mi_rows * 4 + 63) // 64


def partition(sb_rows, world):
    """Contiguous, near-equal split of `sb_rows` superblock rows over `world` ranks: [(first, count), ...].
    Ranks beyond the number of rows get (first, 0) and sit the picture out."""
    out = []
    for g in range(world):
        a, b = g * sb_rows // world, (g + 1) * sb_rows // world
        out.append((a, b - a))
    return out


def stripe_geom(full_geom, first, count):
    """Geometry a context needs (max_geom) to hold stripe (first, count) with its context rows."""
    rows = full_geom.mi_rows * 4
    sb = sb_rows_of(full_geom)
    g0 = 64 * (first - (1 if first > 0 else 0))
    g1 = min(rows, 64 * (first + count + (1 if first + count < sb else 0)))
    return F.make_geom_coded(full_geom.width, min(full_geom.height, g1) - g0, full_geom.mi_cols, (g1 - g0) // 4,
                             full_geom.bit_depth, full_geom.sample_bytes)


def fb_rows(first, count):
    """Filter-block rows (64x64) a stripe owns == its superblock rows."""
    return first, first + count


def gather_search(mse, skip, geom, parts, rank, dist=None, device=None):
    """All-gather of the per-stripe search results.  mse [2][nfb][64] u64 and skip [nfb] u8 are whole-picture arrays
    in which this rank filled its rows.  Returns complete arrays on every rank.  Rows are padded to the largest
    stripe so one all_gather moves everything (8K: 8160 fb x 1 KiB = 8.4 MB in total)."""
    if dist is None or len(parts) == 1:
        return mse, skip
    import torch

    nhfb = (geom.mi_cols + 15) // 16
    maxc = max(c for _, c in parts)
    first, count = parts[rank]
    # int64 view: NCCL has no uint64; the bits travel unchanged
    buf = np.zeros((maxc * nhfb, 2 * 64 + 1), dtype=np.int64)
    sl = slice(first * nhfb, (first + count) * nhfb)
    buf[: count * nhfb, :64] = mse[0, sl].view(np.int64)
    buf[: count * nhfb, 64:128] = mse[1, sl].view(np.int64)
    buf[: count * nhfb, 128] = skip[sl]
    t = torch.from_numpy(buf)
    if device is not None:
        t = t.to(device)
    outs = [torch.empty_like(t) for _ in parts]
    dist.all_gather(outs, t)
    mse_all, skip_all = np.zeros_like(mse), np.zeros_like(skip)
    for (f, c), o in zip(parts, outs):
        o = o.cpu().numpy()
        s2 = slice(f * nhfb, (f + c) * nhfb)
        mse_all[0, s2] = o[: c * nhfb, :64].view(np.uint64)
        mse_all[1, s2] = o[: c * nhfb, 64:128].view(np.uint64)
        skip_all[s2] = o[: c * nhfb, 128].astype(np.uint8)
    return mse_all, skip_all


class StripeFilter:
    """One rank's share of a picture.  Mirrors FilterContext's picture-level calls."""

    def __init__(self, full_geom, rank, world, device=0):
        self.full, self.rank, self.world = full_geom, rank, world
        self.parts = partition(sb_rows_of(full_geom), world)
        self.first, self.count = self.parts[rank]
        self.ctx = None
        if self.count > 0:
            self.ctx = F.FilterContext(stripe_geom(full_geom, self.first, self.count), device=device)

    def close(self):
        if self.ctx is not None:
            self.ctx.close()
            self.ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def begin(self):
        if self.ctx is not None:
            self.ctx.begin_stripe(self.full, self.first, self.count)

    def chain(self, planes, mi, lf, cdef, idx, ccso, flags, out):
        """Upload this rank's rows of `planes`, run deblock -> CDEF(+CCSO), write its rows of `out`."""
        if self.ctx is None:
            return
        c = self.ctx
        c.begin_stripe(self.full, self.first, self.count)
        c.upload(planes)
        c.set_mi(mi)
        c.filter_chain(lf, cdef, idx, ccso, flags)
        c.download(out)

    def search_and_pick(self, rec, src, mi, lf, cfg, dist=None, device=None):
        """Deblock + CDEF strength search on this rank's rows, all-gather, pick (identical on every rank)."""
        g = self.full
        nfb = ((g.mi_rows + 15) // 16) * ((g.mi_cols + 15) // 16)
        mse = np.zeros((2, nfb, 64), dtype=np.uint64)
        skip = np.zeros(nfb, dtype=np.uint8)
        dirs = np.zeros((nfb, 64), dtype=np.uint8)
        var = np.zeros((nfb, 64), dtype=np.int32)
        if self.ctx is not None:
            c = self.ctx
            c.begin_stripe(g, self.first, self.count)
            c.upload(rec)
            c.upload_source(src)
            c.set_mi(mi)
            if lf is not None:
                c.loop_filter_frame(lf)
            c.cdef_seg_search(cfg, out=(mse, skip, dirs, var))
        mse, skip = gather_search(mse, skip, g, self.parts, self.rank, dist, device)
        if self.ctx is not None:
            params, idx = self.ctx.pick_table(cfg, mse, skip)  # GPU pick, identical on every rank
        else:
            params, idx = F.cdef_pick_host(cfg, mse, skip)
        return params, idx, mse, skip
