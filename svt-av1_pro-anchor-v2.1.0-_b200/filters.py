"""Host-side mirror of the reference's frame-level filter interface, over the C ABI.

Method names follow the reference entry points they replace:

  FilterContext.loop_filter_frame   <- svt_av1_loop_filter_frame   (Encoder/Codec/EbDeblockingFilter.c:624)
  FilterContext.cdef_frame          <- svt_av1_cdef_frame          (Encoder/Codec/EbEncCdef.c:284)
  FilterContext.ccso_frame          <- ccso_frame                  (Common/Codec/EbCcso.c:626)
  FilterContext.cdef_seg_search     <- cdef_seg_search             (Encoder/Codec/EbCdefProcess.c:114)
  FilterContext.finish_cdef_search  <- finish_cdef_search          (Encoder/Codec/EbEncCdef.c:728)

Pictures are numpy arrays (uint8 or uint16) in host memory; everything else is plain values.  This
module only marshals; all arithmetic happens in libsvtfilt_b200.so on the GPU (no CPU fallback).
"""
import ctypes as C

import numpy as np

from . import _lib as L


class SvtfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"svtf error {code}: {msg}")
        self.code = code


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def make_geom(width, height, bit_depth, sample_bytes=None):
    """Geometry of a picture: coded size = dimensions rounded up to 8 (mi_cols/mi_rows even)."""
    if sample_bytes is None:
        sample_bytes = 1 if bit_depth == 8 else 2
    return L.Geom(width, height, 2 * ((width + 7) // 8), 2 * ((height + 7) // 8), bit_depth, sample_bytes)


def make_geom_coded(width, height, mi_cols, mi_rows, bit_depth, sample_bytes):
    """Geometry with an explicit coded size (stripes: the coded rows are whole superblock rows)."""
    return L.Geom(width, height, mi_cols, mi_rows, bit_depth, sample_bytes)


def lf_params(filter_level=(0, 0), filter_level_u=0, filter_level_v=0, sharpness_level=0, mode_ref_delta_enabled=0,
              ref_deltas=(1, 0, 0, 0, -1, 0, -1, -1), mode_deltas=(0, 0), segmentation_enabled=0, seg_feature_enabled=None,
              seg_feature_data=None, delta_lf_present=0, delta_lf_multi=0, curr_delta_lf=(0, 0, 0, 0)):
    p = L.LfParams()
    p.filter_level[0], p.filter_level[1] = filter_level
    p.filter_level_u, p.filter_level_v, p.sharpness_level = filter_level_u, filter_level_v, sharpness_level
    p.mode_ref_delta_enabled = mode_ref_delta_enabled
    for i, v in enumerate(ref_deltas):
        p.ref_deltas[i] = v
    for i, v in enumerate(mode_deltas):
        p.mode_deltas[i] = v
    p.segmentation_enabled = segmentation_enabled
    if seg_feature_enabled is not None:
        for s in range(8):
            for f in range(8):
                p.seg_feature_enabled[s][f] = int(seg_feature_enabled[s][f])
                p.seg_feature_data[s][f] = int(seg_feature_data[s][f])
    p.delta_lf_present, p.delta_lf_multi = delta_lf_present, delta_lf_multi
    for i, v in enumerate(curr_delta_lf):
        p.curr_delta_lf[i] = v
    return p


def cdef_params(damping, y_strengths, uv_strengths):
    p = L.CdefParams()
    p.cdef_damping = damping
    n = len(y_strengths)
    p.cdef_bits = max(0, (n - 1).bit_length())
    for i in range(n):
        p.cdef_y_strength[i] = y_strengths[i]
        p.cdef_uv_strength[i] = uv_strengths[i]
    return p


def cdef_search_cfg(first_pass, second_pass, first_pass_uv=None, second_pass_uv=None, subsampling_factor=1, base_q_idx=120,
                    zero_fs_cost_bias=0, lambda_=0):
    c = L.CdefSearchCfg()
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
    p = L.CcsoParams()
    for i in range(3):
        p.ccso_enable[i], p.ccso_bo_only[i], p.quant_idx[i] = enable[i], bo_only[i], quant_idx[i]
        p.ext_filter_support[i], p.edge_clf[i], p.max_band_log2[i] = ext_filter_support[i], edge_clf[i], max_band_log2[i]
    if filter_offset is not None:
        fo = np.ascontiguousarray(filter_offset, dtype=np.int8).reshape(3, L.CCSO_LUT_SIZE)
        C.memmove(p.filter_offset, fo.ctypes.data, 3 * L.CCSO_LUT_SIZE)
    return p


def pack_mi_enc(bsize, tx_depth, skip, ref_frame0, mode, segment_id):
    """svtf_pack_mi_enc over 2-D uint8 arrays of shape (mi_rows, mi_cols) -> uint32 grid."""
    lib = L.load()
    arrs = [np.ascontiguousarray(a, dtype=np.uint8) for a in (bsize, tx_depth, skip, ref_frame0, mode, segment_id)]
    out = np.empty(arrs[0].shape, dtype=np.uint32)
    r = lib.svtf_pack_mi_enc(arrs[0].size, *[_ptr(a) for a in arrs], _ptr(out))
    if r:
        raise SvtfError(r, lib.svtf_last_error(None).decode())
    return out


def pack_mi_dec(bsize, tx_size_y, tx_size_uv, skip, ref_frame0, mode, segment_id):
    lib = L.load()
    arrs = [np.ascontiguousarray(a, dtype=np.uint8) for a in (bsize, tx_size_y, tx_size_uv, skip, ref_frame0, mode, segment_id)]
    out = np.empty(arrs[0].shape, dtype=np.uint32)
    r = lib.svtf_pack_mi_dec(arrs[0].size, *[_ptr(a) for a in arrs], _ptr(out))
    if r:
        raise SvtfError(r, lib.svtf_last_error(None).decode())
    return out


def cdef_pick_host(cfg, mse, skip):
    """svtf_cdef_pick_host: finish_cdef_search on a host table mse[2][nfb][64] (uint64), skip[nfb]."""
    lib = L.load()
    mse = np.ascontiguousarray(mse, dtype=np.uint64)
    skip = np.ascontiguousarray(skip, dtype=np.uint8)
    nfb = skip.size
    out = L.CdefParams()
    idx = np.empty(nfb, dtype=np.int8)
    r = lib.svtf_cdef_pick_host(C.byref(cfg), _ptr(mse), _ptr(skip), nfb, C.byref(out), _ptr(idx))
    if r:
        raise SvtfError(r, lib.svtf_last_error(None).decode())
    return out, idx


def wiener_from_stats(win, M, H):
    """svtf_lr_wiener_from_stats: the host half of search_wiener_seg (Encoder/Codec/EbRestorationPick.c:1296) for one unit.
    Returns (usable, vfilter, hfilter)."""
    lib = L.load()
    M, H = np.ascontiguousarray(M, dtype=np.int64), np.ascontiguousarray(H, dtype=np.int64)
    vf, hf = (C.c_int16 * 8)(), (C.c_int16 * 8)()
    r = lib.svtf_lr_wiener_from_stats(win, _ptr(M), _ptr(H), vf, hf)
    if r < 0:
        raise SvtfError(r, lib.svtf_last_error(None).decode())
    return r, tuple(vf), tuple(hf)


class Comm:
    """svtf_comm: an NCCL communicator owned by the library (svtf_comm_create).  `unique_id()` on one rank, the 128 bytes go to the
    others by any transport (torch.distributed broadcast, a pipe, ...), then every rank constructs Comm(device, nranks, rank, id)."""

    @staticmethod
    def unique_id():
        lib = L.load()
        buf = C.create_string_buffer(128)
        r = lib.svtf_comm_unique_id(buf)
        if r:
            raise SvtfError(r, lib.svtf_last_error(None).decode())
        return buf.raw

    def __init__(self, device, nranks, rank, uid):
        self._lib = L.load()
        h = C.c_void_p()
        r = self._lib.svtf_comm_create(device, nranks, rank, uid, C.byref(h))
        if r:
            raise SvtfError(r, self._lib.svtf_last_error(None).decode())
        self.handle, self.nranks, self.rank = h, nranks, rank

    def close(self):
        if self.handle:
            self._lib.svtf_comm_destroy(self.handle)
            self.handle = None


class PinnedPicture:
    """Three planes in page-locked host memory (svtf_host_alloc), exposed as numpy arrays."""

    def __init__(self, geom):
        lib = L.load()
        self._lib = lib
        dt = np.uint8 if geom.sample_bytes == 1 else np.uint16
        self.planes, self._ptrs = [], []
        for p in range(3):
            w, h = (geom.mi_cols * 4) >> (1 if p else 0), (geom.mi_rows * 4) >> (1 if p else 0)
            nbytes = w * h * geom.sample_bytes
            ptr = lib.svtf_host_alloc(nbytes)
            if not ptr:
                raise SvtfError(L.E_NOMEM, "svtf_host_alloc failed")
            self._ptrs.append(ptr)
            buf = (C.c_uint8 * nbytes).from_address(ptr)
            self.planes.append(np.frombuffer(buf, dtype=dt).reshape(h, w))

    def close(self):
        for ptr in self._ptrs:
            self._lib.svtf_host_free(ptr)
        self._ptrs, self.planes = [], []


class PinnedArray:
    """A numpy array in page-locked host memory (svtf_host_alloc): copies from it are real asynchronous DMA."""

    def __init__(self, shape, dtype):
        lib = L.load()
        self._lib = lib
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._ptr = lib.svtf_host_alloc(nbytes)
        if not self._ptr:
            raise SvtfError(L.E_NOMEM, "svtf_host_alloc failed")
        buf = (C.c_uint8 * nbytes).from_address(self._ptr)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)

    def close(self):
        if self._ptr:
            self._lib.svtf_host_free(self._ptr)
        self._ptr, self.array = None, None


class FilterContext:
    """One picture in flight on one GPU (svtf_ctx)."""

    def __init__(self, max_geom, device=0):
        self._lib = L.load()
        self._h = C.c_void_p()
        self.max_geom = max_geom
        r = self._lib.svtf_ctx_create(device, C.byref(max_geom), C.byref(self._h))
        if r:
            raise SvtfError(r, self._lib.svtf_last_error(None).decode())
        self.geom = None

    def close(self):
        if self._h:
            self._lib.svtf_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _chk(self, r):
        if r:
            raise SvtfError(r, self._lib.svtf_last_error(self._h).decode())

    # ---- picture / side info ----
    def begin(self, geom):
        self.geom = geom
        self._chk(self._lib.svtf_frame_begin(self._h, C.byref(geom)))

    def begin_stripe(self, full_geom, first_sb_row, num_sb_rows):
        """This context filters superblock rows [first, first+num) of a picture (svtf_stripe_begin).  All later calls
        keep taking whole-picture arrays; only the stripe's part of them is read or written."""
        self.geom = full_geom
        self._chk(self._lib.svtf_stripe_begin(self._h, C.byref(full_geom), first_sb_row, num_sb_rows))

    def _planes_args(self, planes):
        args = []
        for a in planes:
            assert a.strides[1] == a.itemsize and a.itemsize == self.geom.sample_bytes
            args += [_ptr(a), a.strides[0] // a.itemsize]
        return args

    def upload(self, planes):
        self._chk(self._lib.svtf_upload_frame(self._h, *self._planes_args(planes)))

    def upload_pre_cdef(self, planes):
        self._chk(self._lib.svtf_upload_pre_cdef(self._h, *self._planes_args(planes)))

    def upload_source(self, planes):
        self._chk(self._lib.svtf_upload_source(self._h, *self._planes_args(planes)))

    def set_mi(self, mi):
        if not (isinstance(mi, np.ndarray) and mi.dtype == np.uint32 and mi.flags.c_contiguous):
            mi = np.ascontiguousarray(mi, dtype=np.uint32)
        self._mi_keep = mi
        self._chk(self._lib.svtf_set_mi(self._h, _ptr(mi), mi.shape[1]))

    def set_mi_soa(self, soa):
        """Side info as byte planes, packed on the device.  soa: dict with bsize, skip, ref_frame0, mode, segment_id and either
        tx_depth (encoder, svtf_set_mi_soa_enc) or tx_size_y + tx_size_uv (decoder, svtf_set_mi_soa_dec)."""
        dec = "tx_size_y" in soa
        keys = ("bsize", "tx_size_y", "tx_size_uv", "skip", "ref_frame0", "mode", "segment_id") if dec else \
            ("bsize", "tx_depth", "skip", "ref_frame0", "mode", "segment_id")
        arrs = [soa[k] if (isinstance(soa[k], np.ndarray) and soa[k].dtype == np.uint8 and soa[k].flags.c_contiguous)
                else np.ascontiguousarray(soa[k], dtype=np.uint8) for k in keys]
        self._mi_keep = arrs
        fn = self._lib.svtf_set_mi_soa_dec if dec else self._lib.svtf_set_mi_soa_enc
        self._chk(fn(self._h, *[_ptr(a) for a in arrs], arrs[0].shape[1]))

    def get_mi(self):
        out = np.zeros((self.geom.mi_rows, self.geom.mi_cols), dtype=np.uint32)
        self._chk(self._lib.svtf_get_mi(self._h, _ptr(out), out.shape[1]))
        return out

    def download(self, out=None):
        g = self.geom
        if out is None:
            dt = np.uint8 if g.sample_bytes == 1 else np.uint16
            out = [np.zeros(((g.mi_rows * 4) >> s, (g.mi_cols * 4) >> s), dtype=dt) for s in (0, 1, 1)]
        self._chk(self._lib.svtf_download_frame(self._h, *self._planes_args(out)))
        return out

    def upload_8as16(self, planes8):
        """svt_convert_pic_8bit_to_16bit on the way in: uint8 planes into a context with 16-bit samples."""
        assert all(p.dtype == np.uint8 and p.flags.c_contiguous for p in planes8)
        args = []
        for a in planes8:
            args += [_ptr(a), a.strides[0]]
        self._chk(self._lib.svtf_upload_frame_8as16(self._h, *args))

    def ref_padded_shapes(self, width, height, org_x, org_y):
        return [((((height + s) >> s) + 2 * ((org_y + s) >> s)), (((width + s) >> s) + 2 * ((org_x + s) >> s))) for s in (0, 1, 1)]

    def download_ref_padded(self, width, height, org_x, org_y, pad_right=0, pad_bottom=0, split=False, two_bit=False, out=None):
        """pad_ref_and_set_flags: the current picture as padded reference planes; with `split` also the 8-bit planes and, with
        `two_bit`, the 2-bit planes of svt_aom_un_pack2d.  Returns (planes, planes8 or None, planesn or None).
        out: optional (planes, planes8, planesn) of preallocated (e.g. pinned) arrays."""
        g = self.geom
        lay = L.RefLayout(width, height, org_x, org_y, pad_right, pad_bottom)
        dt = np.uint8 if g.sample_bytes == 1 else np.uint16
        shp = self.ref_padded_shapes(width, height, org_x, org_y)

        def arr(lst):
            return (C.c_void_p * 3)(*[a.ctypes.data for a in lst]), (C.c_int32 * 3)(*[a.strides[0] // a.itemsize for a in lst])

        if out is not None:
            out, o8, on = out
        else:
            out = [np.zeros(sh, dtype=dt) for sh in shp]
            o8 = [np.zeros(sh, dtype=np.uint8) for sh in shp] if split else None
            on = [np.zeros(sh, dtype=np.uint8) for sh in shp] if two_bit else None
        p, st = arr(out)
        p8, st8 = arr(o8) if split else (None, None)
        pn, stn = arr(on) if two_bit else (None, None)
        self._chk(self._lib.svtf_download_ref_padded(self._h, C.byref(lay), p, st, p8, st8, pn, stn))
        return out, o8, on

    def ref_padded_resident(self, width, height, org_x, org_y, split=True, two_bit=True):
        """The hand-off kernel alone (dst == NULL): padded planes into the device staging block, stream-ordered, no DMA."""
        lay = L.RefLayout(width, height, org_x, org_y, 0, 0)
        one = (C.c_void_p * 3)(1, 1, 1)  # non-NULL markers: which planes to produce
        self._chk(self._lib.svtf_download_ref_padded(self._h, C.byref(lay), None, None, one if split else None, None, one if two_bit else None, None))

    def rewind(self):
        self._chk(self._lib.svtf_rewind(self._h))

    def sync(self):
        self._chk(self._lib.svtf_sync(self._h))

    @property
    def launch_count(self):
        return int(self._lib.svtf_launch_count(self._h))

    @property
    def cuda_stream(self):
        return self._lib.svtf_cuda_stream(self._h)

    def set_stream(self, cuda_stream):
        """Run on a caller-owned cudaStream_t (int handle, e.g. torch.cuda.Stream().cuda_stream); None = own stream."""
        self._chk(self._lib.svtf_ctx_set_stream(self._h, C.c_void_p(cuda_stream) if cuda_stream else None))

    def dev_planes(self, which=1):
        arr = (L.DevPlane * 3)()
        self._chk(self._lib.svtf_dev_planes(self._h, which, arr))
        return [(arr[i].ptr, arr[i].pitch, arr[i].width, arr[i].height) for i in range(3)]

    # ---- filters (reference entry-point names) ----
    def loop_filter_frame(self, lf, plane_start=0, plane_end=3, sb_delta_lf=None, sb_size_log2=6):
        """svt_av1_loop_filter_frame; with sb_delta_lf ([sb_rows, sb_cols, 4] int32) the decoder's per-superblock
        loop-filter deltas replace lf.curr_delta_lf (svtf_deblock_frame_sb_delta)."""
        if sb_delta_lf is None:
            self._chk(self._lib.svtf_deblock_frame(self._h, C.byref(lf), plane_start, plane_end))
            return
        d = np.ascontiguousarray(sb_delta_lf, dtype=np.int32)
        self._chk(self._lib.svtf_deblock_frame_sb_delta(self._h, C.byref(lf), plane_start, plane_end, _ptr(d), d.shape[1], sb_size_log2))

    def deblock_prepare(self):
        """svtf_deblock_prepare: build the edge-code planes of the uploaded mode info now (one launch)."""
        self._chk(self._lib.svtf_deblock_prepare(self._h))

    def cdef_frame(self, cdef, fb_strength_idx):
        idx = np.ascontiguousarray(fb_strength_idx, dtype=np.int8)
        self._chk(self._lib.svtf_cdef_apply_frame(self._h, C.byref(cdef), _ptr(idx)))

    @staticmethod
    def _flag_ptrs(blk_flags):
        if blk_flags is None:
            return None, None
        keep = [None if f is None else np.ascontiguousarray(f, dtype=np.uint8) for f in blk_flags]
        arr = (C.c_void_p * 3)(*[None if k is None else k.ctypes.data for k in keep])
        return arr, keep

    def ccso_frame(self, ccso, blk_flags=None):
        arr, keep = self._flag_ptrs(blk_flags)
        self._chk(self._lib.svtf_ccso_apply_frame(self._h, C.byref(ccso), arr))

    def filter_chain(self, lf=None, cdef=None, fb_strength_idx=None, ccso=None, blk_flags=None):
        idx = None if fb_strength_idx is None else np.ascontiguousarray(fb_strength_idx, dtype=np.int8)
        arr, keep = self._flag_ptrs(blk_flags)
        self._chk(self._lib.svtf_filter_chain(self._h, None if lf is None else C.byref(lf), None if cdef is None else C.byref(cdef),
                                              _ptr(idx), None if ccso is None else C.byref(ccso), arr))

    def cdef_seg_search(self, cfg, want_outputs=True, out=None):
        """out: optional (mse, skip, dirs, var) whole-picture arrays to fill (a striped context fills its rows)."""
        g = self.geom
        nfb = ((g.mi_rows + 15) // 16) * ((g.mi_cols + 15) // 16)
        if not want_outputs:
            self._chk(self._lib.svtf_cdef_search_frame(self._h, C.byref(cfg), None, None, None, None))
            return None
        if out is not None:
            mse, skip, dirs, var = out
        else:
            mse = np.zeros((2, nfb, 64), dtype=np.uint64)
            skip = np.zeros(nfb, dtype=np.uint8)
            dirs = np.zeros((nfb, 64), dtype=np.uint8)
            var = np.zeros((nfb, 64), dtype=np.int32)
        self._chk(self._lib.svtf_cdef_search_frame(self._h, C.byref(cfg), _ptr(mse), _ptr(skip), _ptr(dirs), _ptr(var)))
        return mse, skip, dirs, var

    def loop_restoration_filter_frame(self, lr, units):
        """svt_av1_loop_restoration_filter_frame on the current picture (svtf_lr_apply_frame).  lr: LrParams; units: three
        ctypes arrays of LrUnit (vert_units x horz_units, raster order; None for a plane without restoration)."""
        up = (C.c_void_p * 3)(*[C.cast(u, C.c_void_p) if u is not None else None for u in units])
        self._chk(self._lib.svtf_lr_apply_frame(self._h, C.byref(lr), up))

    def superres_upscale_frame(self, upscaled_width, denominator, filters, tile_col_start_mi=None):
        """svt_av1_superres_upscale_frame (Encoder/Codec/EbRestProcess.c:405) on the current picture (and the picture before CDEF):
        svtf_superres_upscale_frame.  filters: int16 [64][8] = av1_resize_filter_normative.  Afterwards self.geom is the
        upscaled geometry."""
        sr = L.SuperresParams()
        starts = list(tile_col_start_mi) if tile_col_start_mi is not None else [0, self.geom.mi_cols]
        sr.upscaled_width, sr.superres_denominator, sr.tile_cols = upscaled_width, denominator, len(starts) - 1
        for j, v in enumerate(starts):
            sr.tile_col_start_mi[j] = v
        f = np.ascontiguousarray(filters, dtype=np.int16)
        sr.filters = f.ctypes.data
        self._chk(self._lib.svtf_superres_upscale_frame(self._h, C.byref(sr)))
        if denominator != 8:
            self.geom = make_geom_coded(upscaled_width, self.geom.height, 2 * ((upscaled_width + 7) >> 3), self.geom.mi_rows, self.geom.bit_depth,
                                        self.geom.sample_bytes)

    def restoration_search(self, cfg):
        """restoration_seg_search + rest_finish_search (Encoder/Codec/EbRestorationPick.c:1471/1555, Wiener tools) on the
        current picture against the uploaded source (svtf_lr_search_frame).  cfg: LrSearchCfg.  Returns (frame types,
        [LrUnit arrays], [LrUnitSearch arrays]) — the unit arrays are ready for loop_restoration_filter_frame."""
        ftype = (C.c_int32 * 3)()
        units, search = [], []
        for p in range(3):
            hu, vu = self.lr_units(p, cfg.restoration_unit_size[p])
            units.append((L.LrUnit * (hu * vu))())
            search.append((L.LrUnitSearch * (hu * vu))())
        up = (C.c_void_p * 3)(*[C.cast(u, C.c_void_p) for u in units])
        sp = (C.c_void_p * 3)(*[C.cast(u, C.c_void_p) for u in search])
        self.sg_frame_ep_cnt = (C.c_int32 * 16)()
        self._chk(self._lib.svtf_lr_search_frame(self._h, C.byref(cfg), ftype, up, sp, self.sg_frame_ep_cnt))
        return list(ftype), units, search

    def lr_units(self, plane, unit_size):
        hu, vu = C.c_int32(), C.c_int32()
        self._lib.svtf_lr_units(C.byref(self.geom), plane, unit_size, C.byref(hu), C.byref(vu))
        return hu.value, vu.value

    def compute_stats(self, plane, unit_size, wiener_win):
        """svt_av1_compute_stats[_highbd] of every restoration unit of a plane (svtf_lr_wiener_stats): M [units][win^2],
        H [units][win^2][win^2], int64."""
        hu, vu = self.lr_units(plane, unit_size)
        n = wiener_win * wiener_win
        M, Hm = np.zeros((hu * vu, n), np.int64), np.zeros((hu * vu, n, n), np.int64)
        self._chk(self._lib.svtf_lr_wiener_stats(self._h, plane, unit_size, wiener_win, _ptr(M), _ptr(Hm)))
        return M, Hm

    def pick_filter_level(self, start_lf, dlf_avg=0, early_exit_convergence=0, only_4x4=0, search_chroma=1):
        """svtf_lf_search_frame: svt_av1_pick_filter_level(LPF_PICK_FROM_FULL_IMAGE) on the uploaded picture + source."""
        cfg = L.LfSearchCfg(dlf_avg, early_exit_convergence, only_4x4, search_chroma)
        out = (C.c_int32 * 4)()
        self._chk(self._lib.svtf_lf_search_frame(self._h, C.byref(start_lf), C.byref(cfg), out))
        return list(out)

    def ccso_search(self, rdmult, plane_start=0, plane_end=3):
        """svtf_ccso_derive_frame: derive the CCSO parameters and block flags of the current picture against the source."""
        g = self.geom
        out = L.CcsoParams()
        flags = []
        for p in range(3):
            pw, ph, bl = g.width >> (1 if p else 0), g.height >> (1 if p else 0), 7 if p else 8
            flags.append(np.zeros(((pw + (1 << bl) - 1) >> bl) * ((ph + (1 << bl) - 1) >> bl), dtype=np.uint8))
        fl = (C.c_void_p * 3)(*[f.ctypes.data for f in flags])
        self._chk(self._lib.svtf_ccso_derive_frame(self._h, rdmult, plane_start, plane_end, C.byref(out), fl))
        return out, flags

    def pick_table(self, cfg, mse, skip):
        """svtf_cdef_pick_table: finish_cdef_search on the GPU from a host table mse[2][nfb][64], skip[nfb]."""
        mse = np.ascontiguousarray(mse, dtype=np.uint64)
        skip = np.ascontiguousarray(skip, dtype=np.uint8)
        out = L.CdefParams()
        idx = np.empty(skip.size, dtype=np.int8)
        self._chk(self._lib.svtf_cdef_pick_table(self._h, C.byref(cfg), _ptr(mse), _ptr(skip), skip.size, C.byref(out), _ptr(idx)))
        return out, idx

    def upload_own(self, planes):
        """svtf_upload_frame_own: only the rows this (striped) context owns."""
        self._chk(self._lib.svtf_upload_frame_own(self._h, *self._planes_args(planes)))

    def set_mi_own(self, mi):
        mi = np.ascontiguousarray(mi, dtype=np.uint32)
        self._keep_mi = mi
        self._chk(self._lib.svtf_set_mi_own(self._h, _ptr(mi), mi.shape[1]))

    def stripe_exchange(self, comm, rank_up, rank_down):
        """svtf_stripe_exchange: halo rows to / from the neighbouring stripes' GPUs (NCCL send / recv)."""
        self._chk(self._lib.svtf_stripe_exchange(self._h, comm.handle, rank_up, rank_down))

    def search_gather_pick(self, comm, cfg, parts, nfb_full):
        """svtf_cdef_search_gather_pick: device-to-device gather of every rank's mse / skip rows, then the pick."""
        first = np.ascontiguousarray([p[0] for p in parts], dtype=np.int32)
        count = np.ascontiguousarray([p[1] for p in parts], dtype=np.int32)
        out = L.CdefParams()
        idx = np.empty(nfb_full, dtype=np.int8)
        self._chk(self._lib.svtf_cdef_search_gather_pick(self._h, comm.handle, C.byref(cfg), _ptr(first), _ptr(count), C.byref(out), _ptr(idx)))
        return out, idx

    def pick_table_noalloc(self, cfg, mse, skip):
        """svtf_cdef_pick_table_noalloc: the same pick on the library's static device scratch array (no allocation)."""
        mse = np.ascontiguousarray(mse, dtype=np.uint64)
        skip = np.ascontiguousarray(skip, dtype=np.uint8)
        out = L.CdefParams()
        idx = np.empty(skip.size, dtype=np.int8)
        self._chk(self._lib.svtf_cdef_pick_table_noalloc(self._h, C.byref(cfg), _ptr(mse), _ptr(skip), skip.size, C.byref(out), _ptr(idx)))
        return out, idx

    def finish_cdef_search(self, cfg):
        g = self.geom
        nfb = ((g.mi_rows + 15) // 16) * ((g.mi_cols + 15) // 16)
        out = L.CdefParams()
        idx = np.empty(nfb, dtype=np.int8)
        self._chk(self._lib.svtf_cdef_pick(self._h, C.byref(cfg), C.byref(out), _ptr(idx)))
        return out, idx
