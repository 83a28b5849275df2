"""Stand-alone ctypes mirrors of the few include/svtfilt_b200.h parameter structs that the CPU checkers
(oracle/liborc.so, oracle/_ref/libsvtref*.so) take, plus their builders — numpy + ctypes only.

Used INSTEAD of the product package when tests/harness.py is imported with SVTF_HARNESS_NO_PRODUCT=1: `bench.py --impl
reference` must time the reference's CPU code in a process that neither imports the product's Python package nor maps
libsvtfilt_b200.so.  tests/test_host_logic.py pins these layouts against the product binding's."""
import ctypes as C

import numpy as np

CDEF_MAX_STRENGTHS = 16
TOTAL_STRENGTHS = 64
CCSO_LUT_SIZE = 128 * 16


class Geom(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("mi_cols", C.c_int32), ("mi_rows", C.c_int32), ("bit_depth", C.c_int32),
                ("sample_bytes", C.c_int32)]


class LfParams(C.Structure):
    _fields_ = [("filter_level", C.c_int32 * 2), ("filter_level_u", C.c_int32), ("filter_level_v", C.c_int32), ("sharpness_level", C.c_int32),
                ("mode_ref_delta_enabled", C.c_int32), ("ref_deltas", C.c_int8 * 8), ("mode_deltas", C.c_int8 * 2),
                ("segmentation_enabled", C.c_int32), ("seg_feature_enabled", (C.c_int16 * 8) * 8), ("seg_feature_data", (C.c_int16 * 8) * 8),
                ("delta_lf_present", C.c_int32), ("delta_lf_multi", C.c_int32), ("curr_delta_lf", C.c_int32 * 4)]


class CdefParams(C.Structure):
    _fields_ = [("cdef_damping", C.c_int32), ("cdef_bits", C.c_int32), ("cdef_y_strength", C.c_uint8 * CDEF_MAX_STRENGTHS),
                ("cdef_uv_strength", C.c_uint8 * CDEF_MAX_STRENGTHS)]


class CdefSearchCfg(C.Structure):
    _fields_ = [("first_pass_fs_num", C.c_int32), ("default_second_pass_fs_num", C.c_int32),
                ("default_first_pass_fs", C.c_uint8 * TOTAL_STRENGTHS), ("default_second_pass_fs", C.c_uint8 * TOTAL_STRENGTHS),
                ("default_first_pass_fs_uv", C.c_int8 * TOTAL_STRENGTHS), ("default_second_pass_fs_uv", C.c_int8 * TOTAL_STRENGTHS),
                ("subsampling_factor", C.c_int32), ("base_q_idx", C.c_int32), ("zero_fs_cost_bias", C.c_int32), ("lambda_", C.c_uint32)]


class CcsoParams(C.Structure):
    _fields_ = [("ccso_enable", C.c_uint8 * 3), ("ccso_bo_only", C.c_uint8 * 3), ("quant_idx", C.c_uint8 * 3), ("ext_filter_support", C.c_uint8 * 3),
                ("edge_clf", C.c_uint8 * 3), ("max_band_log2", C.c_uint8 * 3), ("filter_offset", (C.c_int8 * CCSO_LUT_SIZE) * 3)]


MIRRORED = ("Geom", "LfParams", "CdefParams", "CdefSearchCfg", "CcsoParams")


def make_geom(width, height, bit_depth, sample_bytes=None):
    if sample_bytes is None:
        sample_bytes = 1 if bit_depth == 8 else 2
    return Geom(width, height, 2 * ((width + 7) // 8), 2 * ((height + 7) // 8), bit_depth, sample_bytes)


def lf_params(filter_level=(0, 0), filter_level_u=0, filter_level_v=0, sharpness_level=0):
    p = LfParams()
    p.filter_level[0], p.filter_level[1] = filter_level
    p.filter_level_u, p.filter_level_v, p.sharpness_level = filter_level_u, filter_level_v, sharpness_level
    for i, v in enumerate((1, 0, 0, 0, -1, 0, -1, -1)):
        p.ref_deltas[i] = v
    return p


def cdef_params(damping, y_strengths, uv_strengths):
    p = CdefParams()
    p.cdef_damping = damping
    n = len(y_strengths)
    p.cdef_bits = max(0, (n - 1).bit_length())
    for i in range(n):
        p.cdef_y_strength[i] = y_strengths[i]
        p.cdef_uv_strength[i] = uv_strengths[i]
    return p


def cdef_search_cfg(first_pass, second_pass, first_pass_uv=None, second_pass_uv=None, subsampling_factor=1, base_q_idx=120, zero_fs_cost_bias=0,
                    lambda_=0):
    c = CdefSearchCfg()
    c.first_pass_fs_num, c.default_second_pass_fs_num = len(first_pass), len(second_pass)
    first_pass_uv = first_pass if first_pass_uv is None else first_pass_uv
    second_pass_uv = second_pass if second_pass_uv is None else second_pass_uv
    for i, v in enumerate(first_pass):
        c.default_first_pass_fs[i] = v
        c.default_first_pass_fs_uv[i] = first_pass_uv[i]
    for i, v in enumerate(second_pass):
        c.default_second_pass_fs[i] = v
        c.default_second_pass_fs_uv[i] = second_pass_uv[i]
    c.subsampling_factor, c.base_q_idx, c.zero_fs_cost_bias, c.lambda_ = subsampling_factor, base_q_idx, zero_fs_cost_bias, lambda_
    return c


def ccso_params(enable=(1, 1, 1), bo_only=(0, 0, 0), quant_idx=(0, 0, 0), ext_filter_support=(0, 0, 0), edge_clf=(0, 0, 0),
                max_band_log2=(0, 0, 0), filter_offset=None):
    p = CcsoParams()
    for i in range(3):
        p.ccso_enable[i], p.ccso_bo_only[i], p.quant_idx[i] = enable[i], bo_only[i], quant_idx[i]
        p.ext_filter_support[i], p.edge_clf[i], p.max_band_log2[i] = ext_filter_support[i], edge_clf[i], max_band_log2[i]
    if filter_offset is not None:
        fo = np.ascontiguousarray(filter_offset, dtype=np.int8).reshape(3, CCSO_LUT_SIZE)
        C.memmove(p.filter_offset, fo.ctypes.data, 3 * CCSO_LUT_SIZE)
    return p


def _no_product(*_a, **_k):
    raise RuntimeError("tests/harness.py was imported with SVTF_HARNESS_NO_PRODUCT=1: the product library is not available in this process")


pack_mi_enc = pack_mi_dec = FilterContext = _no_product
