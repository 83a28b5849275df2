"""Host logic of the single-picture stripe split: partition, stripe geometry and the one real
exchange of the path — the all-gather of per-stripe CDEF-search results before finish_cdef_search — run with
world_size 2 and 3 over gloo on the CPU.  The per-stripe mse here comes from the oracle (the checker), masked to
the rows each rank owns; the gather + pick code is the product's (svtfilt_b200.stripes)."""
import os
import socket

import numpy as np
import pytest


def test_partition_covers_every_row_once(H):
    P = H.S.stripes.partition
    for rows in (1, 2, 3, 7, 17, 34, 68, 135):
        for world in (1, 2, 3, 4, 8, 16):
            parts = P(rows, world)
            assert len(parts) == world
            covered = []
            for f, c in parts:
                assert c >= 0
                covered += list(range(f, f + c))
            assert covered == list(range(rows))
            counts = [c for _, c in parts]
            assert max(counts) - min(counts) <= 1


def test_stripe_geometry(H):
    S = H.S
    g = S.make_geom(7680, 4320, 10)
    sb = S.stripes.sb_rows_of(g)
    assert sb == 68
    for world in (2, 4, 8):
        for f, c in S.stripes.partition(sb, world):
            l = S.stripes.stripe_geom(g, f, c)
            top = 64 if f > 0 else 0
            bot = 64 if f + c < sb else 0
            own = min(4320, 64 * (f + c)) - 64 * f
            assert l.mi_rows * 4 == top + own + bot
            assert l.mi_cols == g.mi_cols and l.width == g.width
    # a picture whose visible height ends inside the last coded 8 rows
    g = S.make_geom(200, 130, 8)  # coded 136 rows, 3 SB rows
    l = S.stripes.stripe_geom(g, 2, 1)
    assert l.mi_rows * 4 == 64 + 8 and l.height == 64 + 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import sys

    import torch.distributed as dist

    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here)
    import harness as H

    S = H.S
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = S.make_geom(200, 264, 10)  # 5 SB rows x 4 fb columns
        rec = H.synth_picture(g, 4100)
        src = H.synth_source(g, rec, 4101)
        mi = H.pack_mi(H.synth_mode_info(g, 4102))
        first, second = H.cdef_level1_lists()
        cfg = S.cdef_search_cfg(first[:8], second[:6], base_q_idx=120, lambda_=3000)
        orc = H.Oracle()
        mse, skip, _, _ = orc.cdef_search(g, rec, src, mi, cfg)
        parts = S.stripes.partition(S.stripes.sb_rows_of(g), world)
        nhfb = (g.mi_cols + 15) // 16
        f, c = parts[rank]
        mine_mse, mine_skip = np.zeros_like(mse), np.zeros_like(skip)
        sl = slice(f * nhfb, (f + c) * nhfb)
        mine_mse[:, sl], mine_skip[sl] = mse[:, sl], skip[sl]
        got_mse, got_skip = S.stripes.gather_search(mine_mse, mine_skip, g, parts, rank, dist)
        ok = np.array_equal(got_mse, mse) and np.array_equal(got_skip, skip)
        a = S.cdef_pick_host(cfg, got_mse, got_skip)
        b = orc.cdef_pick(cfg, mse, skip)
        ok = ok and [a[0].cdef_bits, *a[0].cdef_y_strength, *a[0].cdef_uv_strength] == [b[0].cdef_bits, *b[0].cdef_y_strength, *b[0].cdef_uv_strength]
        ok = ok and np.array_equal(a[1], b[1])
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_search_gather_and_pick_gloo(world):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(r, True) for r in range(world)]
