"""GPU: the one-launch strength pick (pick_cluster_kernel, svtf_pick2.cuh: a 16-CTA cluster resident for all 75 svt_search_one_dual
passes; selected with SVTF_PICK_CLUSTER=1, read once per process, hence the subprocess) against the single-thread host implementation:
ties (first index wins), zero-strength bias, skipped blocks, short lists, a wide table (falls back to the per-pass launches)."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = textwrap.dedent("""
    import sys, os
    sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
    import numpy as np
    import harness as H
    S = H.S
    rng = np.random.default_rng(11)
    first, second = H.cdef_level1_lists()
    sig = lambda p: [p.cdef_bits, p.cdef_damping, *p.cdef_y_strength, *p.cdef_uv_strength]
    with S.FilterContext(S.make_geom(64, 64, 8)) as ctx:
        for it, (nfb, n1, n2, bias, pskip, hi) in enumerate(((1, 16, 48, 0, 0.0, 22), (7, 3, 3, 0, 0.2, 22), (60, 16, 48, 40, 0.2, 22), (2040, 16, 48, 0, 0.1, 22),
                                                            (500, 16, 20, 70, 1.0, 22), (4080, 16, 48, 0, 0.3, 30), (300, 16, 48, 0, 0.1, 40),
                                                            (8160, 16, 48, 0, 0.1, 22), (8160, 11, 30, 30, 0.2, 28))):
            cfg = S.cdef_search_cfg(first[:n1], second[:n2], base_q_idx=int(rng.integers(0, 256)), zero_fs_cost_bias=bias, lambda_=int(rng.integers(1, 50000)))
            mse = rng.integers(0, 1 << hi, (2, nfb, 64)).astype(np.uint64)
            if it == 3:
                mse = (mse >> 12) << 12
            skip = (rng.random(nfb) < pskip).astype(np.uint8)
            l0 = ctx.launch_count
            a = ctx.pick_table(cfg, mse, skip)
            b = S.cdef_pick_host(cfg, mse, skip)
            assert sig(a[0]) == sig(b[0]) and np.array_equal(a[1], b[1]), it
            print("case", it, "launches", ctx.launch_count - l0)
    print("ok")
""") % (ROOT, ROOT)


def test_cluster_pick_matches_host_pick():
    env = dict(os.environ, SVTF_PICK_CLUSTER="1", SVTF_PICK_GRID="0")
    r = subprocess.run([sys.executable, "-c", SCRIPT], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
    launches = {int(l.split()[1]): int(l.split()[3]) for l in r.stdout.splitlines() if l.startswith("case")}
    assert launches[3] == 1, launches          # narrow table: one launch
    assert launches[6] > 70, launches          # entries >= 2^31: the cluster kernel declines, the per-pass launches run


def test_grid_pick_matches_host_pick():
    """The cooperative whole-device form (pick_grid_kernel, svtf_pick3.cuh; the default) on the same cases."""
    env = dict(os.environ)
    env.pop("SVTF_PICK_CLUSTER", None), env.pop("SVTF_PICK_GRID", None)
    r = subprocess.run([sys.executable, "-c", SCRIPT], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
    launches = {int(l.split()[1]): int(l.split()[3]) for l in r.stdout.splitlines() if l.startswith("case")}
    assert launches[3] == 1 and launches[5] == 1, launches   # narrow tables that fit the device: one launch
    assert launches[7] == 1 and launches[8] == 1, launches   # 8K: two row groups per CTA (32-bit and 64-bit partial sums)
    assert launches[6] > 70, launches                        # entries >= 2^31: declined, the per-pass launches run


def test_per_pass_pick_matches_host_pick():
    """The per-pass form (pick_tot_kernel, svtf_pick.cuh; what 8K tables and wide tables run) forced on the same cases."""
    env = dict(os.environ, SVTF_PICK_GRID="0")
    env.pop("SVTF_PICK_CLUSTER", None)
    r = subprocess.run([sys.executable, "-c", SCRIPT], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
    launches = {int(l.split()[1]): int(l.split()[3]) for l in r.stdout.splitlines() if l.startswith("case")}
    assert min(launches.values()) > 70, launches
