"""CPU: the C-ABI library loads, exports every symbol include/svtfilt_b200.h declares, and fails
loudly (no CPU fallback) when there is no CUDA device.  No compute calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch


def _declared_symbols(H):
    txt = open(os.path.join(H.ROOT, "include", "svtfilt_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(svtf_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(H):
    lib = H.L.load()
    declared = _declared_symbols(H)
    assert len(declared) >= 25
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, f"declared in the header but not exported: {missing}"
    assert sorted(H.L.EXPORTS) == declared, "python binding list out of sync with the header"


def test_struct_sizes_match_the_header(H, tmp_path):
    """ctypes mirrors vs the C compiler's view of include/svtfilt_b200.h."""
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "svtfilt_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(svtf_geom),sizeof(svtf_lf_params),sizeof(svtf_cdef_params),sizeof(svtf_cdef_search_cfg),"
                   "sizeof(svtf_ccso_params),sizeof(svtf_dev_plane));return 0;}\n")
    exe = tmp_path / "sz"
    import subprocess

    subprocess.run(["gcc", "-I", os.path.join(H.ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(v) for v in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    L = H.L
    assert got == [C.sizeof(L.Geom), C.sizeof(L.LfParams), C.sizeof(L.CdefParams), C.sizeof(L.CdefSearchCfg), C.sizeof(L.CcsoParams),
                   C.sizeof(L.DevPlane)]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU error path")
def test_no_gpu_means_error_not_fallback(H):
    S = H.S
    with pytest.raises(S.SvtfError) as e:
        S.FilterContext(S.make_geom(64, 64, 8))
    assert e.value.code == H.L.E_CUDA
    assert "no CPU fallback" in str(e.value)


def test_bad_geometry_rejected(H):
    S = H.S
    lib = H.L.load()
    h = C.c_void_p()
    for g in (H.L.Geom(0, 64, 16, 16, 8, 1), H.L.Geom(64, 64, 15, 16, 8, 1), H.L.Geom(64, 64, 16, 16, 9, 1),
              H.L.Geom(64, 64, 16, 16, 10, 1), H.L.Geom(128, 64, 16, 16, 8, 1)):
        assert lib.svtf_ctx_create(0, C.byref(g), C.byref(h)) == H.L.E_ARG


def test_pack_rejects_bad_enums(H):
    S = H.S
    one = np.zeros((1, 1), np.uint8)
    with pytest.raises(S.SvtfError):
        S.pack_mi_enc(one + 22, one, one, one, one, one)
    with pytest.raises(S.SvtfError):
        S.pack_mi_dec(one, one + 19, one, one, one, one, one)


def test_null_context_calls_return_errors(H):
    lib = H.L.load()
    assert lib.svtf_sync(None) == H.L.E_ARG
    assert lib.svtf_rewind(None) == H.L.E_ARG
    assert lib.svtf_deblock_frame(None, None, 0, 3) == H.L.E_ARG
    assert b"NULL" in lib.svtf_last_error(None)
