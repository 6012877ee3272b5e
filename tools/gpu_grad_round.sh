#!/bin/bash
# Gradient visit: parity tests, full-vector check against the prototype, gradient timing with and without the sub-chunk pipeline.
set -u
TAG=${1:-r01y}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest.log
timeout 300 python tools/gpu_grad_check.py > $OUT/${TAG}_gradcheck.log 2>&1; echo "gradcheck exit $?"; tail -12 $OUT/${TAG}_gradcheck.log
timeout 300 python - > $OUT/${TAG}_gradtime.log 2>&1 <<'P'
import sys, numpy as np, torch
sys.path.insert(0, ".")
from physics_informed_reinforcement_learning_b200 import RCWABatch
for order in (40, 60):
    B = 4096
    big = torch.from_numpy(np.random.default_rng(1234).integers(0, 2, (B, 256)).astype(np.int8) * 2 - 1).cuda()
    be = RCWABatch(256, order, max_batch=B, device=0)
    ref = None
    for sub in (0, -1):
        be.set_subchunk(sub)
        for _ in range(3): out = be.evaluate_grad(big)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): out = be.evaluate_grad(big)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        same = None if ref is None else (torch.equal(out[0], ref[0]), torch.equal(out[1], ref[1]))
        ref = out
        print(f"order {order} subchunk {sub}: {ms:.2f} ms / {B} -> {B / ms * 1e3:.0f} grads/s  bit-equal to serial: {same}")
P
cat $OUT/${TAG}_gradtime.log
timeout 600 python bench.py --steps 10 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"; cat $OUT/${TAG}_bench.json
