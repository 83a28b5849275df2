#!/usr/bin/env bash
# Builds the reference applications twice, out of tree, from /root/reference (read-only):
#   integration/_bin/SvtAv1EncApp_ref,  SvtAv1DecApp_ref   — the unmodified reference (C only: there is no nasm/yasm in this image)
#   integration/_bin/SvtAv1EncApp_b200, SvtAv1DecApp_b200  — the same tree + integration/svt_b200.patch (-DSVT_B200_FILTERS=ON): the
#       filter stages call adapters/svtf_*_adapters.c over libsvtfilt_b200.so
# The patched tree is a COPY under integration/_build/src_b200 (git- and gpurun-ignored); the binaries travel to the GPU box.
# Needs: cmake, ninja, gcc, patch; the reference tree at $REF; the CUDA library already built (__graft_entry__.build()).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="$(dirname "$HERE")"
REF="${REF:-/root/reference}"
B="$HERE/_build"
mkdir -p "$B" "$HERE/_bin"
COMMON=(-G Ninja -DCMAKE_BUILD_TYPE=Release -DCOMPILE_C_ONLY=ON -DBUILD_SHARED_LIBS=OFF -DBUILD_TESTING=OFF)

if [ ! -x "$HERE/_bin/SvtAv1EncApp_ref" ] || [ "${FORCE:-0}" = 1 ]; then
  cmake "${COMMON[@]}" -S "$REF" -B "$B/ref" -DCMAKE_OUTPUT_DIRECTORY="$B/ref/bin/" > "$B/ref_cmake.log" 2>&1
  ninja -C "$B/ref" > "$B/ref_ninja.log" 2>&1
  cp "$B/ref/bin/SvtAv1EncApp" "$HERE/_bin/SvtAv1EncApp_ref"
  cp "$B/ref/bin/SvtAv1DecApp" "$HERE/_bin/SvtAv1DecApp_ref"
fi

# patched copy: only the directories the build needs; re-created only when the patch changed (keeps the ninja build incremental)
if [ ! -f "$B/src_b200/.patched" ] || [ "$HERE/svt_b200.patch" -nt "$B/src_b200/.patched" ]; then
  rm -rf "$B/src_b200"
  mkdir -p "$B/src_b200"
  for t in CMakeLists.txt Source cmake third_party Config; do cp -rp "$REF/$t" "$B/src_b200/"; done
  patch -s -p1 -d "$B/src_b200" < "$HERE/svt_b200.patch"
  touch "$B/src_b200/.patched"
fi
cmake "${COMMON[@]}" -S "$B/src_b200" -B "$B/b200" -DSVT_B200_FILTERS=ON -DSVTFILT_DIR="$ROOT" -DCMAKE_OUTPUT_DIRECTORY="$B/b200/bin/" > "$B/b200_cmake.log" 2>&1
ninja -C "$B/b200" > "$B/b200_ninja.log" 2>&1
cp "$B/b200/bin/SvtAv1EncApp" "$HERE/_bin/SvtAv1EncApp_b200"
cp "$B/b200/bin/SvtAv1DecApp" "$HERE/_bin/SvtAv1DecApp_b200"
# the decoder's multi-threaded path is not reachable from SvtAv1DecApp (-threads is disabled there): a minimal API client
LIBDIR="$ROOT/svt-av1_pro-anchor-v2.1.0-_b200/lib"
gcc -O2 -I"$REF/Source/API" -o "$HERE/_bin/dec_threads_ref" "$HERE/dec_threads_main.c" "$B/ref/bin/libSvtAv1Dec.a" -lpthread -lm
gcc -O2 -I"$REF/Source/API" -o "$HERE/_bin/dec_threads_b200" "$HERE/dec_threads_main.c" "$B/b200/bin/libSvtAv1Dec.a" \
    "$LIBDIR/libsvtfilt_b200.so" -Wl,-rpath,"$LIBDIR" -lpthread -lm
ls -la "$HERE/_bin"
