#!/usr/bin/env bash
# Patched vs reference encoder over a matrix of encoder options: md5 of the reconstruction and of the bitstream.
set -u
B=integration/_bin; T=$(mktemp -d); cd "$(dirname "$0")/.."
mk() { python - <<PY
import sys; sys.path.insert(0,'integration'); import synth_yuv; synth_yuv.write_clip('$T/c_$1x$2_$3.yuv',$1,$2,$4,$3)
PY
}
run() { # name w h depth frames preset extra...
  local name=$1 W=$2 H=$3 D=$4 N=$5 P=$6; shift 6
  [ -f $T/c_${W}x${H}_${D}.yuv ] || mk $W $H $D 6
  for app in ref b200; do
    SVTF_ADAPTER_STATS=1 timeout 900 $B/SvtAv1EncApp_$app -i $T/c_${W}x${H}_${D}.yuv -w $W -h $H --fps 30 -n $N --preset $P --input-depth $D "$@" -b $T/$app.ivf -o $T/$app.yuv > $T/$app.log 2>&1 || echo "  enc $app FAILED rc=$? $(tail -2 $T/$app.log | tr '\n' ' ')"
  done
  local a=$(md5sum < $T/ref.yuv) b=$(md5sum < $T/b200.yuv) c=$(md5sum < $T/ref.ivf) d=$(md5sum < $T/b200.ivf)
  if [ "$a" = "$b" ] && [ "$c" = "$d" ]; then echo "OK        $name"; else echo "MISMATCH  $name (recon $([ "$a" = "$b" ] && echo same || echo DIFF), bitstream $([ "$c" = "$d" ] && echo same || echo DIFF))"; fi
}
if [ "${SET:-1}" = 2 ]; then
for P in 3 5 7 9 11 13; do run "preset $P 8-bit" 640 360 8 5 $P --crf 40; done
for P in 3 5 9 13; do run "preset $P 10-bit" 416 240 10 5 $P --crf 40; done
run "fast-decode 1"             640 360 8 5 6 --fast-decode 1 --crf 40
run "tune 0 (VQ)"               640 360 8 5 6 --tune 0 --crf 40
run "tune 2 (SSIM)"             640 360 8 5 6 --tune 2 --crf 40
run "lp 1"                      640 360 8 5 6 --lp 1 --crf 40
run "keyint 2"                  640 360 8 6 6 --keyint 2 --crf 40
run "crf 20 (weak filters)"     640 360 8 4 6 --crf 20
run "crf 63 (strong filters)"   640 360 8 4 6 --crf 63
run "1280x720 preset 4 10-bit"  1280 720 10 3 4 --crf 45
run "1920x1080 preset 5"        1920 1080 8 3 5 --crf 45
run "enable-dlf 2"              640 360 8 4 6 --enable-dlf 2 --crf 45
run "cdef off"                  640 360 8 4 6 --enable-cdef 0 --crf 45
run "restoration off"           640 360 8 4 4 --enable-restoration 0 --crf 45
run "vbr 1-pass"                640 360 8 6 8 --rc 1 --tbr 800
exit 0
fi
run "aq-mode 1 (segmentation)"      640 360 8 5 6 --aq-mode 1 --crf 45
run "aq-mode 2 (delta q)"           640 360 8 5 6 --aq-mode 2 --crf 45
run "2x2 tiles"                     640 360 8 5 6 --tile-rows 1 --tile-columns 1 --crf 45
run "2 tile cols + superres"        640 360 8 4 6 --tile-columns 1 --superres-mode 1 --superres-denom 11 --superres-kf-denom 11 --crf 50
run "odd size 638x358"              638 358 8 4 6 --crf 45
run "odd size 250x150 10-bit"       250 150 10 4 4 --crf 45
run "resize-mode 1 denom 12"        640 360 8 4 6 --resize-mode 1 --resize-denom 12 --resize-kf-denom 12 --crf 45
run "superres random (mode 2)"      640 360 8 5 6 --superres-mode 2 --crf 50
run "low-delay, preset 10"          640 360 8 6 10 --pred-struct 1 --crf 45
run "sharpness 3"                   640 360 8 4 6 --sharpness 3 --crf 45
run "preset 0 tiny"                 192 128 8 2 0 --crf 45
run "preset 12"                     640 360 8 6 12 --crf 45
run "screen content"                640 360 8 3 6 --scm 1 --crf 45
