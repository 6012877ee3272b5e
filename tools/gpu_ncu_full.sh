#!/bin/bash
# ncu --set full capture of selected kernels of the bench command (run only after the plain command exited 0).
# usage: bash tools/gpu_ncu_full.sh <tag> <kernel-regex> [batch]
set -u
TAG=$1; KRE=$2; BATCH=${3:-1024}
OUT=gpurun_out; mkdir -p $OUT
BCMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-grad --batch $BATCH"
timeout 300 $BCMD > $OUT/${TAG}_plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 3 -c 5 -o $OUT/${TAG}_prof -f $BCMD > $OUT/${TAG}_ncufull.log 2>&1
echo "ncu full exit $?"; tail -3 $OUT/${TAG}_ncufull.log
