#!/bin/bash
# k_apply at the headline size (order 40, batch 4096): sweeps per systolic pass 4 (default) / 6 / 8
for k in 0 6 8; do
  echo "PIRL_APPLY81=$k"
  PIRL_APPLY81=$k timeout 200 python - <<'P'
import sys; sys.path.insert(0, ".")
import numpy as np, torch
from physics_informed_reinforcement_learning_b200 import RCWABatch
B = 4096
be = RCWABatch(256, 40, max_batch=B)
pats = torch.from_numpy((np.random.default_rng(5).integers(0, 2, (B, 256)) * 2 - 1).astype(np.int8)).cuda()
e0 = be.evaluate(pats); torch.cuda.synchronize()
be.set_subchunk(0); be.set_stage_timing(True); be.stage_times()
for _ in range(5): e = be.evaluate(pats)
torch.cuda.synchronize()
t = be.stage_times()
be.set_stage_timing(False); be.set_subchunk(-1)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for _ in range(10): e = be.evaluate(pats)
ev[1].record(); torch.cuda.synchronize()
print({k: round(v[0] / 5, 3) for k, v in t.items() if v[1]}, "pipelined ms/step %.3f" % (ev[0].elapsed_time(ev[1]) / 10), "checksum", float(e.sum()))
P
done
