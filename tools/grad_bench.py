"""Times pirl_rcwa_eval_grad on the bench workload (4096 random 256-cell patterns, order 40); also the ncu target for the
gradient kernels:  ncu --set full -k regex:"g_tail|g_eigback" -s 2 -c 2 python tools/grad_bench.py"""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from physics_informed_reinforcement_learning_b200 import RCWABatch
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ORDER = int(sys.argv[2]) if len(sys.argv) > 2 else 40
big = torch.from_numpy(np.random.default_rng(1234).integers(0, 2, (B, 256)).astype(np.int8) * 2 - 1).cuda()
be = RCWABatch(256, ORDER, max_batch=B, device=0)
for _ in range(2): be.evaluate_grad(big)
torch.cuda.synchronize()
be.set_stage_timing(True); be.stage_times()
t0 = time.time()
for _ in range(3): be.evaluate_grad(big)
torch.cuda.synchronize()
dt = (time.time() - t0) / 3
print(f"eval_grad {B}: {dt * 1e3:.2f} ms -> {B / dt:.0f} grads/s", {k: round(v[0] / 3, 3) for k, v in be.stage_times().items() if v[1]})
