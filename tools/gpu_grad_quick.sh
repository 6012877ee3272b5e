#!/bin/bash
# Quick gradient visit: gradient parity tests, per-phase clocks, gradient timing at orders 40 / 60.
set -u
TAG=${1:-r01z}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_grad.py -m gpu -x -q > $OUT/${TAG}_pytest_grad.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest_grad.log
timeout 300 python tools/gpu_grad_check.py > $OUT/${TAG}_gradcheck.log 2>&1; echo "gradcheck exit $?"; grep -E "N=|host" $OUT/${TAG}_gradcheck.log
timeout 300 python tools/phase_clocks.py 4096 grad > $OUT/${TAG}_phase_grad.txt 2>&1; cat $OUT/${TAG}_phase_grad.txt
timeout 300 python tools/grad_bench.py 4096 40 2>&1 | tail -1 | tee $OUT/${TAG}_gradbench.txt
timeout 300 python tools/grad_bench.py 4096 60 2>&1 | tail -1 | tee -a $OUT/${TAG}_gradbench.txt
