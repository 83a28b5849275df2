"""The decoder adapters are 4:2:0 only (out of scope: the reference decoder also reads 4:0:0 / 4:2:2 / 4:4:4,
Decoder/Codec/EbDecCdef.c:172-173).  A stream of another format must stop the process with a message — never be filtered as if it
were 4:2:0 — and it must do so before any device call, so this runs without a GPU."""
import os, subprocess, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "adapters", "_build", "libsvtadapt_dec_test.so")


@pytest.mark.skipif(not os.path.exists(LIB), reason="adapter harness not built (needs the reference headers)")
@pytest.mark.parametrize("mono,ssx,ssy", [(1, 1, 1), (0, 0, 0), (0, 1, 0)])
def test_other_chroma_formats_stop_loudly(mono, ssx, ssy):
    code = f"import ctypes; l = ctypes.CDLL({LIB!r}); l.harness_dec_format_probe({mono}, {ssx}, {ssy}); print('RETURNED')"
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "RETURNED" not in r.stdout, (r.returncode, r.stdout, r.stderr[-400:])
    assert "only 4:2:0 streams are supported" in r.stderr, r.stderr[-400:]
