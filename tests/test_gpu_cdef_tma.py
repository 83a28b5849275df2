"""GPU: the TMA-fed persistent variant of the CDEF apply kernel (cdef4_kernel, svtf_cdef4.cuh; selected with SVTF_CDEF_TMA=1, read once
per process, hence the subprocess) produces the oracle's bytes — 8-bit in byte and in 16-bit storage, 10- and 12-bit, odd sizes, with and
without the fused CCSO epilogue."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("w,h,bd,sb", [(328, 200, 8, 1), (328, 200, 10, 2), (136, 72, 12, 2), (72, 264, 8, 2), (1920, 1080, 8, 1), (3840, 2160, 10, 2)])
def test_cdef_tma_variant_matches_oracle(w, h, bd, sb):
    env = dict(os.environ, SVTF_CDEF_TMA="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "c4_probe.py"), str(w), str(h), str(bd), str(sb)], capture_output=True, text=True, env=env,
                       timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert f"cdef {w} {h} {bd} {sb} equal" in r.stdout and "cdef+ccso equal" in r.stdout, r.stdout
