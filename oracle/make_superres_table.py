"""Writes tests/golden/superres_filters.npy: the 64 x 8 filter table of the normative super-resolution upscale
(av1_resize_filter_normative, Common/Codec/EbSuperRes.h:22 — constants of the AV1 specification), read out of the reference
compiled into oracle/_ref/libsvtref.so (ref_superres_filters, oracle/ref_shim_handoff.c).  The library under test embeds no copy:
callers hand the table in (svtf_superres_params.filters); tests and bench.py use this fixture.  Run where /root/reference is
mounted: python oracle/make_superres_table.py"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libsvtref.so"))
lib.ref_superres_filters.restype = C.POINTER(C.c_int16)
tab = np.ctypeslib.as_array(lib.ref_superres_filters(), shape=(64, 8)).copy()
assert (tab.sum(axis=1) == 128).all()
np.save(os.path.join(ROOT, "tests", "golden", "superres_filters.npy"), tab)
print("wrote", tab.shape)
