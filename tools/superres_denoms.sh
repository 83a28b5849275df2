#!/usr/bin/env bash
set -u
B=integration/_bin; T=$(mktemp -d); cd "$(dirname "$0")/.."
W=${W:-640}; H=${H:-360}; N=${N:-2}
python - <<PY
import sys; sys.path.insert(0,'integration'); import synth_yuv; synth_yuv.write_clip('$T/c.yuv',$W,$H,$N,8)
PY
for d in 9 10 11 12 13 14 15 16; do
 for extra in "" "--enable-cdef 0 --enable-restoration 0" ; do
  for app in ref b200; do
    timeout 600 $B/SvtAv1EncApp_$app -i $T/c.yuv -w $W -h $H --fps 30 -n $N --preset 6 --input-depth 8 --crf 50 --superres-mode 1 --superres-denom $d --superres-kf-denom $d ${EXTRA:-} $extra -b $T/$app.ivf -o $T/$app.yuv > $T/$app.log 2>&1 || echo "enc $app failed"
  done
  python - <<PY
import numpy as np
w,h,n=$W,$H,$N
a=np.fromfile('$T/ref.yuv',dtype=np.uint8); b=np.fromfile('$T/b200.yuv',dtype=np.uint8)
fs=w*h*3//2; msg=[]
for f in range(min(len(a),len(b))//fs):
    fa=a[f*fs:(f+1)*fs]; fb=b[f*fs:(f+1)*fs]
    for pn,(o,pw,ph) in enumerate(((0,w,h),(w*h,w//2,h//2),(w*h*5//4,w//2,h//2))):
        pa=fa[o:o+pw*ph].reshape(ph,pw).astype(int); pb=fb[o:o+pw*ph].reshape(ph,pw).astype(int)
        d=np.argwhere(pa!=pb)
        if len(d): msg.append(f"f{f}p{pn}:{len(d)} cols {d[:,1].min()}..{d[:,1].max()}")
print("denom $d narrow", (w*8+$d//2)//$d, "[$extra]:", "IDENTICAL" if not msg else " ".join(msg[:5]))
PY
 done
done
