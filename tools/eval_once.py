"""Minimal driver for ncu: a few forward evaluations of one configuration.  python tools/eval_once.py n_cells order batch [reps] [algo]"""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
from physics_informed_reinforcement_learning_b200 import RCWABatch
n_cells, order, B = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 4
algo = sys.argv[5] if len(sys.argv) > 5 else "TMM"
be = RCWABatch(n_cells, order, max_batch=B, algo=algo)
be.set_subchunk(0)
pats = torch.from_numpy((np.random.default_rng(5).integers(0, 2, (B, n_cells)) * 2 - 1).astype(np.int8)).cuda()
for _ in range(reps):
    eff = be.evaluate(pats)
torch.cuda.synchronize()
print("ok", float(eff.mean()), be.launch_count())
