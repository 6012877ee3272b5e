#!/bin/bash
# k_apply variants at order 100 (PIRL_APPLY = 0 old 64 rows x 2 sweeps, 1: 32x4, 2: 64x4, 3: 32x2): stage times at batch 512
for v in ${VARIANTS:-0 1 2 3 4 5}; do
  echo "PIRL_APPLY=$v"
  PIRL_APPLY=$v timeout 200 python - <<'P'
import sys; sys.path.insert(0, ".")
import numpy as np, torch
from physics_informed_reinforcement_learning_b200 import RCWABatch
B = 512
be = RCWABatch(512, 100, max_batch=B)
pats = torch.from_numpy((np.random.default_rng(5).integers(0, 2, (B, 512)) * 2 - 1).astype(np.int8)).cuda()
e0 = be.evaluate(pats); torch.cuda.synchronize()
be.set_stage_timing(True); be.stage_times()
for _ in range(3): e = be.evaluate(pats)
torch.cuda.synchronize()
t = be.stage_times()
print({k: round(v[0] / 3, 3) for k, v in t.items() if v[1]}, "checksum", float(e.sum()))
P
done
