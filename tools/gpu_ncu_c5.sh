#!/bin/bash
# ncu captures of the cluster path (order 100 / 512 cells, batch 512): launch list + --set full of the four kernels.
# Run only after the plain command exited 0 (same call, && directly before ncu).
set -u
TAG=${1:-r02k}
OUT=gpurun_out; mkdir -p $OUT
CMD="python tools/eval_once.py 512 100 512 4"
timeout 300 $CMD > $OUT/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ck_front|ck_tail2|k_apply|k_ql" -s 12 -c 4 -o $OUT/${TAG}_c5_prof -f $CMD > $OUT/${TAG}_ncufull.log 2>&1
echo "ncu full exit $?"; tail -3 $OUT/${TAG}_ncufull.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 15 -c 5 --csv --log-file $OUT/${TAG}_c5_launches.csv $CMD > $OUT/${TAG}_ncul.log 2>&1
echo "ncu launches exit $?"; tail -8 $OUT/${TAG}_c5_launches.csv
ls -la $OUT/${TAG}*.ncu-rep
