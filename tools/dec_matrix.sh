#!/usr/bin/env bash
# Streams from the reference encoder over a matrix of options, decoded by both decoders (and the multi-threaded API client).
set -u
B=integration/_bin; T=$(mktemp -d); cd "$(dirname "$0")/.."
mk() { python - <<PY
import sys; sys.path.insert(0,'integration'); import synth_yuv; synth_yuv.write_clip('$T/c_$1x$2_$3.yuv',$1,$2,$4,$3)
PY
}
run() { # name w h depth frames preset extra...
  local name=$1 W=$2 H=$3 D=$4 N=$5 P=$6; shift 6
  [ -f $T/c_${W}x${H}_${D}.yuv ] || mk $W $H $D 8
  timeout 900 $B/SvtAv1EncApp_ref -i $T/c_${W}x${H}_${D}.yuv -w $W -h $H --fps 30 -n $N --preset $P --input-depth $D "$@" -b $T/s.ivf -o $T/s.yuv > $T/enc.log 2>&1 || { echo "ENC FAILED $name"; return; }
  timeout 300 $B/SvtAv1DecApp_ref -i $T/s.ivf -o $T/d_ref.yuv > /dev/null 2>&1
  SVTF_ADAPTER_STATS=1 timeout 300 $B/SvtAv1DecApp_b200 -i $T/s.ivf -o $T/d_b200.yuv > $T/dec.log 2>&1 || echo "  dec b200 FAILED: $(grep -a 'Svt\[error\]' $T/dec.log | head -1)"
  timeout 300 $B/dec_threads_b200 $T/s.ivf $T/d_mt.yuv 3 > $T/decmt.log 2>&1 || echo "  dec_threads b200 FAILED"
  local a=$(md5sum < $T/d_ref.yuv) b=$(md5sum < $T/d_b200.yuv) c=$(md5sum < $T/d_mt.yuv)
  echo "$([ "$a" = "$b" ] && echo OK || echo MISMATCH) / mt $([ "$a" = "$c" ] && echo OK || echo MISMATCH)  $name   [$(grep -ao 'dec_loop_filter_frame=.*' $T/dec.log)]"
}
run "aq-mode 1"                     640 360 8 5 6 --aq-mode 1 --crf 45
run "aq-mode 2"                     640 360 8 5 6 --aq-mode 2 --crf 45
run "2x2 tiles"                     640 360 8 5 6 --tile-rows 1 --tile-columns 1 --crf 45
run "tile cols + superres 8/11"     640 360 8 4 6 --tile-columns 1 --superres-mode 1 --superres-denom 11 --superres-kf-denom 11 --crf 50
run "superres 8/9"                  640 360 8 3 6 --superres-mode 1 --superres-denom 9 --superres-kf-denom 9 --crf 50
run "superres random"               640 360 8 6 6 --superres-mode 2 --crf 50
run "resize-mode 1 8/12"            640 360 8 4 6 --resize-mode 1 --resize-denom 12 --resize-kf-denom 12 --crf 45
run "resize-mode 3 (random)"        640 360 8 6 6 --resize-mode 3 --crf 45
run "250x150 10-bit preset 4"       250 150 10 4 4 --crf 45
run "low delay preset 10"           640 360 8 6 10 --pred-struct 1 --crf 45
run "128x128 superblocks (preset 1)" 512 384 8 3 1 --crf 45
run "10-bit superres 8/13 preset 4" 640 360 10 3 4 --superres-mode 1 --superres-denom 13 --superres-kf-denom 13 --crf 50
run "film grain"                    640 360 8 3 6 --film-grain 10 --crf 45
