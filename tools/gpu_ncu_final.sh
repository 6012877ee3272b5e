#!/bin/bash
# ncu --set full captures of the current build: the four forward kernels (bench command, batch 4096, pipeline off) and the two
# gradient kernels (tools/grad_bench.py).  Run only after the plain commands exited 0.
set -u
TAG=${1:-r01zc}
OUT=gpurun_out; mkdir -p $OUT
export PIRL_SUBCHUNK=0
BCMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-grad --batch 4096"
timeout 300 $BCMD > $OUT/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"rk_front|k_ql|k_apply|rk_tail2" -s 12 -c 4 -o $OUT/${TAG}_prof -f $BCMD > $OUT/${TAG}_ncufull.log 2>&1
echo "ncu forward exit $?"; tail -2 $OUT/${TAG}_ncufull.log
GCMD="python tools/grad_bench.py 4096 40"
timeout 300 $GCMD > $OUT/${TAG}_grad_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"g_tail|g_eigback" -s 2 -c 2 -o $OUT/${TAG}_grad_prof -f $GCMD > $OUT/${TAG}_grad_ncufull.log 2>&1
echo "ncu grad exit $?"; tail -2 $OUT/${TAG}_grad_ncufull.log
ls -la $OUT/${TAG}*.ncu-rep
