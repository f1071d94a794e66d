#!/bin/bash
# sweep of the NTT tuning knobs (FB_NTT_MAXR / FB_NTT_TILE / FB_NTT_MAXB / FB_NTT_G / FB_NTT_TDIV); one line per configuration
mkdir -p gpurun_out
out=gpurun_out/ntt_sweep.txt
: > $out
for g in ${GS:-3 2}; do for td in ${TDIVS:-8 4}; do
for cfg in ${CONFIGS:-"10 11 3" "10 10 3" "8 11 3" "8 10 3" "7 11 4"}; do
  set -- $cfg
  echo -n "G=$g TDIV=$td MAXR=$1 TILE=$2 MAXB=$3: " >> $out
  FB_NTT_G=$g FB_NTT_TDIV=$td FB_NTT_MAXR=$1 FB_NTT_TILE=$2 FB_NTT_MAXB=$3 timeout 200 python tools/mini_ntt.py ${SIZES:-16 18 20 22 24} >> $out 2>gpurun_out/ntt_sweep.err || echo "failed" >> $out
done; done; done
cat $out
