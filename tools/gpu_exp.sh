#!/bin/bash
mkdir -p gpurun_out
out=gpurun_out/exp2.txt; : > $out
for cfg in "3 11 4" "4 10 4" "4 11 4" "4 10 8" "4 9 4" "4 9 2"; do
  set -- $cfg
  echo -n "ntt MINB=$1 TILE=$2 TDIV=$3: " >> $out; FB_NTT_MINB=$1 FB_NTT_TILE=$2 FB_NTT_TDIV=$3 timeout 200 python tools/mini_ntt.py 16 18 20 22 24 >> $out 2>&1
done
cat $out
