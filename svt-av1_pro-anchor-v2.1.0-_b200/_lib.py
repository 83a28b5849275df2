"""ctypes binding of libsvtfilt_b200.so (the Tier-1 C ABI, include/svtfilt_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  There is no fallback:
if the shared object is missing, importing this module raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libsvtfilt_b200.so")

MAX_SEGMENTS = 8
SEG_LVL_MAX = 8
REF_FRAMES = 8
CDEF_MAX_STRENGTHS = 16
TOTAL_STRENGTHS = 64
CCSO_LUT_SIZE = 128 * 16

OK, E_ARG, E_CUDA, E_STATE, E_NOMEM = 0, -1, -2, -3, -4


class Geom(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("mi_cols", C.c_int32), ("mi_rows", C.c_int32),
                ("bit_depth", C.c_int32), ("sample_bytes", C.c_int32)]


class LfParams(C.Structure):
    _fields_ = [("filter_level", C.c_int32 * 2), ("filter_level_u", C.c_int32), ("filter_level_v", C.c_int32),
                ("sharpness_level", C.c_int32), ("mode_ref_delta_enabled", C.c_int32),
                ("ref_deltas", C.c_int8 * REF_FRAMES), ("mode_deltas", C.c_int8 * 2),
                ("segmentation_enabled", C.c_int32),
                ("seg_feature_enabled", (C.c_int16 * SEG_LVL_MAX) * MAX_SEGMENTS),
                ("seg_feature_data", (C.c_int16 * SEG_LVL_MAX) * MAX_SEGMENTS),
                ("delta_lf_present", C.c_int32), ("delta_lf_multi", C.c_int32), ("curr_delta_lf", C.c_int32 * 4)]


class CdefParams(C.Structure):
    _fields_ = [("cdef_damping", C.c_int32), ("cdef_bits", C.c_int32),
                ("cdef_y_strength", C.c_uint8 * CDEF_MAX_STRENGTHS), ("cdef_uv_strength", C.c_uint8 * CDEF_MAX_STRENGTHS)]


class CdefSearchCfg(C.Structure):
    _fields_ = [("first_pass_fs_num", C.c_int32), ("default_second_pass_fs_num", C.c_int32),
                ("default_first_pass_fs", C.c_uint8 * TOTAL_STRENGTHS), ("default_second_pass_fs", C.c_uint8 * TOTAL_STRENGTHS),
                ("default_first_pass_fs_uv", C.c_int8 * TOTAL_STRENGTHS), ("default_second_pass_fs_uv", C.c_int8 * TOTAL_STRENGTHS),
                ("subsampling_factor", C.c_int32), ("base_q_idx", C.c_int32), ("zero_fs_cost_bias", C.c_int32),
                ("lambda_", C.c_uint32)]


class CcsoParams(C.Structure):
    _fields_ = [("ccso_enable", C.c_uint8 * 3), ("ccso_bo_only", C.c_uint8 * 3), ("quant_idx", C.c_uint8 * 3),
                ("ext_filter_support", C.c_uint8 * 3), ("edge_clf", C.c_uint8 * 3), ("max_band_log2", C.c_uint8 * 3),
                ("filter_offset", (C.c_int8 * CCSO_LUT_SIZE) * 3)]


class LrUnit(C.Structure):
    """svtf_lr_unit (RestorationUnitInfo, Common/Codec/EbRestoration.h:184-188)"""
    _fields_ = [("restoration_type", C.c_int32), ("wiener_vfilter", C.c_int16 * 8), ("wiener_hfilter", C.c_int16 * 8), ("sgr_ep", C.c_int32),
                ("sgr_xqd", C.c_int32 * 2)]


class LrParams(C.Structure):
    """svtf_lr_params"""
    _fields_ = [("frame_restoration_type", C.c_int32 * 3), ("restoration_unit_size", C.c_int32 * 3), ("optimized_lr", C.c_int32)]


class LrSearchCfg(C.Structure):
    """svtf_lr_search_cfg (WnFilterCtrls / SgFilterCtrls, Common/Codec/Av1Common.h:32-60; Macroblock.rdmult / *_restore_cost)"""
    _fields_ = [("restoration_unit_size", C.c_int32 * 3), ("wn_enabled", C.c_int32), ("wn_filter_tap_lvl", C.c_int32),
                ("wn_use_refinement", C.c_int32), ("wn_max_one_refinement_step", C.c_int32), ("wn_use_chroma", C.c_int32),
                ("sg_enabled", C.c_int32), ("sg_use_chroma", C.c_int32), ("sg_start_ep", C.c_int32 * 2), ("sg_end_ep", C.c_int32 * 2),
                ("sg_ep_inc", C.c_int32 * 2), ("sg_refine", C.c_int32 * 2), ("rdmult", C.c_int32), ("wiener_restore_cost", C.c_int32 * 2),
                ("sgrproj_restore_cost", C.c_int32 * 2), ("switchable_restore_cost", C.c_int32 * 3), ("prev_units", C.c_void_p * 3)]


class LrUnitSearch(C.Structure):
    """svtf_lr_unit_search (RestUnitSearchInfo, Common/Codec/EbRestoration.h:349)"""
    _fields_ = [("sse", C.c_int64 * 3), ("wiener_vfilter", C.c_int16 * 8), ("wiener_hfilter", C.c_int16 * 8), ("sgr_ep", C.c_int32),
                ("sgr_xqd", C.c_int32 * 2), ("reserved", C.c_int32)]


class SuperresParams(C.Structure):
    """svtf_superres_params"""
    _fields_ = [("upscaled_width", C.c_int32), ("superres_denominator", C.c_int32), ("tile_cols", C.c_int32), ("tile_col_start_mi", C.c_int32 * 65),
                ("filters", C.c_void_p)]


class LfSearchCfg(C.Structure):
    _fields_ = [("dlf_avg", C.c_int32), ("early_exit_convergence", C.c_int32), ("only_4x4", C.c_int32), ("search_chroma", C.c_int32)]


class RefLayout(C.Structure):
    """svtf_ref_layout"""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("org_x", C.c_int32), ("org_y", C.c_int32), ("pad_right", C.c_int32),
                ("pad_bottom", C.c_int32)]


class DevPlane(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("pitch", C.c_int32), ("width", C.c_int32), ("height", C.c_int32)]


# every symbol include/svtfilt_b200.h declares (tests check the library exports all of them)
EXPORTS = [
    "svtf_pack_mi_enc", "svtf_pack_mi_dec", "svtf_set_mi_soa_enc", "svtf_set_mi_soa_dec", "svtf_get_mi", "svtf_upload_pre_cdef", "svtf_lr_units", "svtf_lr_apply_frame", "svtf_lf_build_tables", "svtf_ctx_create", "svtf_ctx_destroy", "svtf_last_error", "svtf_sync",
    "svtf_launch_count", "svtf_frame_begin", "svtf_stripe_begin", "svtf_upload_frame", "svtf_upload_source", "svtf_overwrite_current", "svtf_set_mi",
    "svtf_download_frame", "svtf_download_frame_async", "svtf_rewind", "svtf_deblock_frame", "svtf_deblock_prepare", "svtf_deblock_frame_sb_delta", "svtf_cdef_apply_frame", "svtf_ccso_apply_frame",
    "svtf_filter_chain", "svtf_lf_search_frame", "svtf_ccso_derive_frame", "svtf_cdef_search_frame", "svtf_cdef_pick", "svtf_cdef_pick_table", "svtf_cdef_pick_host", "svtf_dev_planes",
    "svtf_cuda_stream", "svtf_ctx_set_stream", "svtf_host_alloc", "svtf_host_free", "svtf_host_register", "svtf_host_unregister", "svtf_download_ref_padded", "svtf_upload_frame_8as16",
    "svtf_cdef_pick_table_noalloc", "svtf_comm_unique_id", "svtf_comm_create", "svtf_comm_destroy", "svtf_upload_frame_own", "svtf_set_mi_own", "svtf_stripe_exchange", "svtf_cdef_search_gather_pick", "svtf_lr_search_frame", "svtf_lr_wiener_stats", "svtf_lr_wiener_from_stats", "svtf_superres_upscale_frame",
]

_lib = None


def load():
    """Load the CUDA library; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(the filter path is CUDA-only, there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, u8p, i8p = C.c_void_p, C.c_int32, C.POINTER(C.c_uint8), C.POINTER(C.c_int8)
    lib.svtf_pack_mi_enc.argtypes = [C.c_size_t, vp, vp, vp, vp, vp, vp, vp]
    lib.svtf_pack_mi_dec.argtypes = [C.c_size_t, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.svtf_lf_build_tables.argtypes = [C.POINTER(LfParams), vp, vp, vp, vp]
    lib.svtf_ctx_create.argtypes = [C.c_int, C.POINTER(Geom), C.POINTER(vp)]
    lib.svtf_ctx_destroy.argtypes = [vp]
    lib.svtf_ctx_destroy.restype = None
    lib.svtf_last_error.argtypes = [vp]
    lib.svtf_last_error.restype = C.c_char_p
    lib.svtf_sync.argtypes = [vp]
    lib.svtf_launch_count.argtypes = [vp]
    lib.svtf_launch_count.restype = C.c_uint64
    lib.svtf_frame_begin.argtypes = [vp, C.POINTER(Geom)]
    lib.svtf_stripe_begin.argtypes = [vp, C.POINTER(Geom), C.c_int32, C.c_int32]
    for f in (lib.svtf_upload_frame, lib.svtf_upload_source, lib.svtf_upload_pre_cdef, lib.svtf_overwrite_current, lib.svtf_download_frame, lib.svtf_download_frame_async):
        f.argtypes = [vp, vp, i32, vp, i32, vp, i32]
    lib.svtf_upload_frame_8as16.argtypes = [vp, vp, i32, vp, i32, vp, i32]
    lib.svtf_download_ref_padded.argtypes = [vp, C.POINTER(RefLayout), C.POINTER(vp), C.POINTER(i32), C.POINTER(vp), C.POINTER(i32), C.POINTER(vp),
                                             C.POINTER(i32)]
    lib.svtf_set_mi.argtypes = [vp, vp, i32]
    lib.svtf_rewind.argtypes = [vp]
    lib.svtf_set_mi_soa_enc.argtypes = [vp] + [vp] * 6 + [i32]
    lib.svtf_set_mi_soa_dec.argtypes = [vp] + [vp] * 7 + [i32]
    lib.svtf_get_mi.argtypes = [vp, vp, i32]
    lib.svtf_lr_units.argtypes = [C.POINTER(Geom), i32, i32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.svtf_lr_apply_frame.argtypes = [vp, C.POINTER(LrParams), C.POINTER(C.c_void_p)]
    lib.svtf_lr_search_frame.argtypes = [vp, C.POINTER(LrSearchCfg), C.POINTER(C.c_int32), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                         C.POINTER(C.c_int32)]
    lib.svtf_superres_upscale_frame.argtypes = [vp, C.POINTER(SuperresParams)]
    lib.svtf_lr_wiener_stats.argtypes = [vp, i32, i32, i32, vp, vp]
    lib.svtf_lr_wiener_from_stats.argtypes = [i32, vp, vp, C.POINTER(C.c_int16), C.POINTER(C.c_int16)]
    lib.svtf_deblock_frame.argtypes = [vp, C.POINTER(LfParams), i32, i32]
    lib.svtf_deblock_prepare.argtypes = [vp]
    lib.svtf_deblock_frame_sb_delta.argtypes = [vp, C.POINTER(LfParams), i32, i32, vp, i32, i32]
    lib.svtf_cdef_apply_frame.argtypes = [vp, C.POINTER(CdefParams), vp]
    lib.svtf_ccso_apply_frame.argtypes = [vp, C.POINTER(CcsoParams), C.POINTER(vp)]
    lib.svtf_filter_chain.argtypes = [vp, C.POINTER(LfParams), C.POINTER(CdefParams), vp, C.POINTER(CcsoParams), C.POINTER(vp)]
    lib.svtf_cdef_search_frame.argtypes = [vp, C.POINTER(CdefSearchCfg), vp, vp, vp, vp]
    lib.svtf_lf_search_frame.argtypes = [vp, C.POINTER(LfParams), C.POINTER(LfSearchCfg), vp]
    lib.svtf_ccso_derive_frame.argtypes = [vp, i32, i32, i32, C.POINTER(CcsoParams), vp]
    lib.svtf_cdef_pick.argtypes = [vp, C.POINTER(CdefSearchCfg), C.POINTER(CdefParams), vp]
    lib.svtf_cdef_pick_table.argtypes = [vp, C.POINTER(CdefSearchCfg), vp, vp, i32, C.POINTER(CdefParams), vp]
    lib.svtf_cdef_pick_table_noalloc.argtypes = [vp, C.POINTER(CdefSearchCfg), vp, vp, i32, C.POINTER(CdefParams), vp]
    lib.svtf_comm_unique_id.argtypes = [vp]
    lib.svtf_comm_create.argtypes = [C.c_int, C.c_int, C.c_int, vp, C.POINTER(vp)]
    lib.svtf_comm_destroy.argtypes = [vp]
    lib.svtf_comm_destroy.restype = None
    lib.svtf_upload_frame_own.argtypes = [vp, vp, i32, vp, i32, vp, i32]
    lib.svtf_set_mi_own.argtypes = [vp, vp, i32]
    lib.svtf_stripe_exchange.argtypes = [vp, vp, i32, i32]
    lib.svtf_cdef_search_gather_pick.argtypes = [vp, vp, C.POINTER(CdefSearchCfg), vp, vp, C.POINTER(CdefParams), vp]
    lib.svtf_cdef_pick_host.argtypes = [C.POINTER(CdefSearchCfg), vp, vp, i32, C.POINTER(CdefParams), vp]
    lib.svtf_dev_planes.argtypes = [vp, C.c_int, C.POINTER(DevPlane)]
    lib.svtf_cuda_stream.argtypes = [vp]
    lib.svtf_cuda_stream.restype = vp
    lib.svtf_ctx_set_stream.argtypes = [vp, vp]
    lib.svtf_host_alloc.argtypes = [C.c_size_t]
    lib.svtf_host_alloc.restype = vp
    lib.svtf_host_register.argtypes = [vp, C.c_size_t]
    lib.svtf_host_unregister.argtypes = [vp]
    lib.svtf_host_free.argtypes = [vp]
    lib.svtf_host_free.restype = None
    _lib = lib
    return lib
