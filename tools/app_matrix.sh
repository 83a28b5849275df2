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
