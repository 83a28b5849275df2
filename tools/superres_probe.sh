#!/usr/bin/env bash
# Bisect: patched vs reference encoder with super-resolution on; per-frame / per-plane first mismatch of the -o reconstruction.
set -u
B=integration/_bin; T=$(mktemp -d); cd "$(dirname "$0")/.."
W=${W:-640}; H=${H:-360}; N=${N:-3}
python - <<PY
import sys; sys.path.insert(0,'integration'); import synth_yuv; synth_yuv.write_clip('$T/c.yuv',$W,$H,$N,8)
PY
run() { # name, extra args
  local name=$1; shift
  for app in ref b200; do
    SVTF_ADAPTER_STATS=1 timeout 600 $B/SvtAv1EncApp_$app -i $T/c.yuv -w $W -h $H --fps 30 -n $N --preset ${PRESET:-6} --input-depth 8 --crf 50 ${EXTRA:-} "$@" -b $T/$app.ivf -o $T/$app.yuv > $T/$app.log 2>&1 || echo "enc $app failed"
  done
  python - <<PY
import numpy as np
w,h,n=$W,$H,$N
a=np.fromfile('$T/ref.yuv',dtype=np.uint8); b=np.fromfile('$T/b200.yuv',dtype=np.uint8)
fs=w*h*3//2
msg=[]
for f in range(min(len(a),len(b))//fs):
    fa=a[f*fs:(f+1)*fs]; fb=b[f*fs:(f+1)*fs]
    for pn,(o,pw,ph) in enumerate(((0,w,h),(w*h,w//2,h//2),(w*h*5//4,w//2,h//2))):
        pa=fa[o:o+pw*ph].reshape(ph,pw).astype(int); pb=fb[o:o+pw*ph].reshape(ph,pw).astype(int)
        d=np.argwhere(pa!=pb)
        if len(d): msg.append(f"f{f} p{pn}: {len(d)} diffs, first (y,x)={tuple(d[0])}, max|d|={abs(pa-pb).max()}, rows {d[:,0].min()}..{d[:,0].max()} cols {d[:,1].min()}..{d[:,1].max()}")
print("$name:", "IDENTICAL" if not msg and len(a)==len(b) else ("sizes %d %d; "%(len(a),len(b)))+" | ".join(msg[:4]))
PY
  grep -o "loop_restoration_filter_frame=[0-9]* restoration_search=[0-9]* superres_upscale_frame=[0-9]*" $T/b200.log
}
SR=${SR:-"--superres-mode 1 --superres-denom 12 --superres-kf-denom 12"}
run "no superres"
run "superres" $SR
run "superres, no restoration" $SR --enable-restoration 0
run "superres, no restoration, no cdef" $SR --enable-restoration 0 --enable-cdef 0
run "superres, no cdef" $SR --enable-cdef 0
# decoders on the reference encoder's super-resolution stream
for app in ref b200; do
  $B/SvtAv1EncApp_ref -i $T/c.yuv -w $W -h $H --fps 30 -n $N --preset ${PRESET:-6} --input-depth 8 --crf 50 $SR -b $T/s.ivf -o $T/s.yuv > /dev/null 2>&1
  SVTF_ADAPTER_STATS=1 timeout 300 $B/SvtAv1DecApp_$app -i $T/s.ivf -o $T/dec_$app.yuv > $T/dec_$app.log 2>&1; echo "dec $app rc=$?"; grep "adapter stats" $T/dec_$app.log
  SVTF_ADAPTER_STATS=1 timeout 300 $B/dec_threads_$app $T/s.ivf $T/decmt_$app.yuv 3 > $T/decmt_$app.log 2>&1; echo "dec_threads $app rc=$?"
done
md5sum $T/s.yuv $T/dec_*.yuv $T/decmt_*.yuv
