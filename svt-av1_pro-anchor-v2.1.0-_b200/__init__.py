"""svtfilt-b200: B200-native in-loop filters (deblock, CDEF, CCSO, CDEF strength search) for the
SVT-AV1 "pro anchor v2.1.0" fork.  The product is the C-ABI library built from csrc/ (see
include/svtfilt_b200.h); this package is the thin Python host used by tests and bench.py.

The directory name is not a Python identifier; import it as ``svtfilt_b200`` (the loader module at
the repository root registers this package under that name).
"""
from . import _lib, stripes
from .filters import (Comm, FilterContext, PinnedArray, PinnedPicture, SvtfError, ccso_params, cdef_params, cdef_pick_host, cdef_search_cfg,
                      lf_params, make_geom, make_geom_coded, pack_mi_dec, pack_mi_enc, wiener_from_stats)
from .stripes import StripeFilter

__all__ = ["Comm", "FilterContext", "PinnedArray", "PinnedPicture", "SvtfError", "ccso_params", "cdef_params", "cdef_pick_host", "cdef_search_cfg",
           "lf_params", "make_geom", "make_geom_coded", "pack_mi_dec", "pack_mi_enc", "wiener_from_stats", "StripeFilter", "stripes", "_lib"]
