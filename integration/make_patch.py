#!/usr/bin/env python
"""Regenerates integration/svt_b200.patch: the edits a maintainer makes to the reference tree so that the encoder / decoder
libraries call the B200 filter path (adapters/svtf_*_adapters.c over libsvtfilt_b200.so) instead of their CPU bodies.

Works on a COPY of the reference (integration/_build/src_b200, git- and gpurun-ignored; never committed): copies the tree,
applies the edits below programmatically, and writes `diff -ruN reference copy` to integration/svt_b200.patch — the committed
artefact.  integration/build.sh applies that patch with patch(1); this script is only needed when the edits change.

The edits (all under `#if SVT_B200_FILTERS`, so the patched tree still builds the stock binaries with the option off):
  * CMake: option SVT_B200_FILTERS + SVTFILT_DIR; the adapters are compiled into SvtAv1Enc / SvtAv1Dec, which link libsvtfilt_b200.so;
  * the CPU bodies the adapters replace are compiled out (one `#if !SVT_B200_FILTERS` per function);
  * three call sites: the CDEF segment search (one whole-picture launch by segment 0), the per-picture context release, and
    resident mode at encoder / decoder init.
"""
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("REF", "/root/reference")
DST = os.path.join(HERE, "_build", "src_b200")
TOP = ("CMakeLists.txt", "Source", "cmake", "third_party", "Config")


def read(rel):
    with open(os.path.join(DST, rel)) as f:
        return f.read()


def write(rel, s):
    with open(os.path.join(DST, rel), "w") as f:
        f.write(s)


def guard_function(src, signature_start, macro="SVT_B200_FILTERS", note=""):
    """Wraps the definition that starts with `signature_start` at the beginning of a line (through its closing brace at column 0)
    in `#if !macro ... #endif`."""
    m = re.search(r"^" + re.escape(signature_start), src, flags=re.M)
    if not m:
        raise SystemExit(f"make_patch: `{signature_start}` not found")
    start = m.start()
    e = re.compile(r"\n\}[ \t]*(\n|$)").search(src, start)
    if not e:
        raise SystemExit(f"make_patch: end of `{signature_start}` not found")
    end = e.end()
    if not src[:end].endswith("\n"):
        src, end = src[:end] + "\n" + src[end:], end + 1
    head = f"#if !{macro} /* replaced by the B200 filter path{note} */\n"
    return src[:start] + head + src[start:end] + f"#endif /* !{macro} */\n" + src[end:]


def replace_once(src, old, new):
    if src.count(old) != 1:
        raise SystemExit(f"make_patch: expected exactly one occurrence of:\n{old}\n(found {src.count(old)})")
    return src.replace(old, new)


def main():
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    os.makedirs(DST)
    for t in TOP:
        s = os.path.join(REF, t)
        (shutil.copytree if os.path.isdir(s) else shutil.copy2)(s, os.path.join(DST, t))

    # ---- CMake ------------------------------------------------------------------------------------------------------
    s = read("CMakeLists.txt")
    s = replace_once(s, 'option(COMPILE_C_ONLY "Compile only C code with no simds (autodetect, default off for x86)" OFF)\n',
                     'option(COMPILE_C_ONLY "Compile only C code with no simds (autodetect, default off for x86)" OFF)\n'
                     'option(SVT_B200_FILTERS "Run DLF / CDEF / CCSO / loop restoration on a B200 through libsvtfilt_b200" OFF)\n'
                     'set(SVTFILT_DIR "" CACHE PATH "Root of the svtfilt_b200 repository (include/, adapters/, the built library)")\n'
                     'if(SVT_B200_FILTERS)\n'
                     '    add_compile_definitions(SVT_B200_FILTERS=1)\n'
                     '    include_directories(${SVTFILT_DIR}/include)\n'
                     '    set(SVTFILT_LIB "${SVTFILT_DIR}/svt-av1_pro-anchor-v2.1.0-_b200/lib/libsvtfilt_b200.so")\n'
                     'endif()\n')
    write("CMakeLists.txt", s)
    for lib, adapter in (("Encoder", "svtf_enc_adapters.c"), ("Decoder", "svtf_dec_adapters.c")):
        rel = f"Source/Lib/{lib}/CMakeLists.txt"
        tgt = "SvtAv1Enc" if lib == "Encoder" else "SvtAv1Dec"
        s = read(rel)
        s = replace_once(s, f"set_target_properties({tgt} PROPERTIES VERSION ",
                         "if(SVT_B200_FILTERS)\n"
                         f"    target_sources({tgt} PRIVATE ${{SVTFILT_DIR}}/adapters/{adapter})\n"
                         f"    target_link_libraries({tgt} PUBLIC ${{SVTFILT_LIB}})\n"
                         "endif()\n"
                         f"set_target_properties({tgt} PROPERTIES VERSION ")
        write(rel, s)

    # ---- encoder: bodies the adapters replace -------------------------------------------------------------------------
    rel = "Source/Lib/Encoder/Codec/EbDeblockingFilter.c"
    s = read(rel)
    for sig in ("void svt_av1_loop_filter_init(PictureControlSet *pcs) {", "void svt_aom_loop_filter_sb(EbPictureBufferDesc *frame_buffer,",
                "void svt_av1_loop_filter_frame(EbPictureBufferDesc *frame_buffer,", "EbErrorType svt_av1_pick_filter_level(EbPictureBufferDesc *srcBuffer,"):
        s = guard_function(s, sig)
    write(rel, s)
    rel = "Source/Lib/Encoder/Codec/EbEncCdef.c"
    s = read(rel)
    for sig in ("void svt_av1_cdef_frame(SequenceControlSet *scs, PictureControlSet *pcs) {", "void finish_cdef_search(PictureControlSet *pcs) {"):
        s = guard_function(s, sig)
    write(rel, s)
    rel = "Source/Lib/Common/Codec/EbCcso.c"
    s = read(rel)
    for sig in ("void extend_ccso_border(uint16_t *buf, const int d, MacroblockdPlane *pd) {", "void ccso_frame(EbPictureBufferDesc *frame,"):
        s = guard_function(s, sig, note=" (encoder builds only: the decoder library does not use these two)")
    write(rel, s)
    rel = "Source/Lib/Encoder/Codec/EbPickccso.c"
    write(rel, guard_function(read(rel), "void ccso_search(PictureControlSet *pcs,"))
    rel = "Source/Lib/Common/Codec/EbRestoration.c"
    write(rel, guard_function(read(rel), "void svt_av1_loop_restoration_filter_frame(int32_t *rst_tmpbuf,",
                              note=" (encoder adapter; the decoder library has its own frame function and never calls this one)"))
    rel = "Source/Lib/Encoder/Codec/EbRestorationPick.c"
    s = read(rel)
    for sig in ("void restoration_seg_search(int32_t *rst_tmpbuf,", "void rest_finish_search(PictureControlSet *pcs) {"):
        s = guard_function(s, sig)
    write(rel, s)
    rel = "Source/Lib/Encoder/Codec/EbRestProcess.c"
    write(rel, guard_function(read(rel), "void svt_av1_superres_upscale_frame(struct Av1Common *cm,"))

    # ---- encoder: call sites ----------------------------------------------------------------------------------------------
    rel = "Source/Lib/Encoder/Codec/EbCdefProcess.c"
    s = read(rel)
    s = replace_once(s, "                cdef_seg_search(pcs, scs, dlf_results->segment_index);\n",
                     "#if SVT_B200_FILTERS /* one whole-picture launch; the other segments only take part in the join below */\n"
                     "                if (dlf_results->segment_index == 0)\n"
                     "                    svtf_cdef_search_picture(pcs, scs);\n"
                     "#else\n"
                     "                cdef_seg_search(pcs, scs, dlf_results->segment_index);\n"
                     "#endif\n")
    s = replace_once(s, '#include "EbPickccso.h"\n', '#include "EbPickccso.h"\n#if SVT_B200_FILTERS\n#include "svtfilt_b200_adapters.h"\n#endif\n')
    write(rel, s)

    # ---- decoder ---------------------------------------------------------------------------------------------------------
    rel = "Source/Lib/Decoder/Codec/EbDecLF.c"
    write(rel, guard_function(read(rel), "void svt_aom_dec_av1_loop_filter_frame(EbDecHandle *dec_handle_ptr,"))
    rel = "Source/Lib/Decoder/Codec/EbDecCdef.c"
    write(rel, guard_function(read(rel), "void svt_cdef_frame(EbDecHandle *dec_handle, int enable_flag) {"))
    rel = "Source/Lib/Decoder/Codec/EbDecCcso.c"
    write(rel, guard_function(read(rel), "void dec_ccso_frame(EbPictureBufferDesc *frame,"))
    rel = "Source/Lib/Decoder/Codec/EbDecRestoration.c"
    s = read(rel)
    for sig in ("void svt_aom_dec_av1_loop_restoration_filter_frame(EbDecHandle *dec_handle,", "void svt_aom_dec_av1_loop_restoration_save_boundary_lines(EbDecHandle *dec_handle,"):
        s = guard_function(s, sig)
    write(rel, s)

    rel = "Source/Lib/Decoder/Codec/EbDecSuperRes.c"
    write(rel, guard_function(read(rel), "void svt_av1_superres_upscale(Av1Common *cm, FrameHeader *frm_hdr, SeqHeader *seq_hdr,"))
    # the multi-threaded variants (every thread calls them; the adapters run the frame-level launch on the calling thread)
    rel = "Source/Lib/Decoder/Codec/EbDecProcess.c"
    s = read(rel)
    for sig in ("void svt_aom_dec_av1_loop_filter_frame_mt(EbDecHandle *dec_handle, EbPictureBufferDesc *recon_picture_buf,",
                "void svt_cdef_frame_mt(EbDecHandle *dec_handle_ptr, DecThreadCtxt *thread_ctxt) {",
                "void svt_aom_dec_av1_loop_restoration_filter_frame_mt(EbDecHandle *dec_handle, DecThreadCtxt *thread_ctxt) {"):
        s = guard_function(s, sig)
    write(rel, s)

    out = os.path.join(HERE, "svt_b200.patch")
    # file by file over the copy (the copy holds only the directories the build needs)
    txt = subprocess.run(["bash", "-c", f"cd {DST} && for f in $(find . -type f | sort); do diff -uN --label a/${{f#./}} --label b/${{f#./}} {REF}/$f $f; done"],
                         capture_output=True, text=True).stdout
    with open(out, "w") as f:
        f.write(txt)
    print(f"wrote {out}: {txt.count(chr(10))} lines, {txt.count('+++ b/')} files")


if __name__ == "__main__":
    sys.exit(main())
