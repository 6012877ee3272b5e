"""Per-phase clock64 breakdown of the register-tile kernels (PIRL_PHASE_CLOCKS=1).  Run on the GPU box."""
import os, sys
os.environ["PIRL_PHASE_CLOCKS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from physics_informed_reinforcement_learning_b200 import RCWABatch
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ORDER = int(sys.argv[3]) if len(sys.argv) > 3 else 40
be = RCWABatch(256, ORDER, max_batch=B)
pats = torch.from_numpy((np.random.default_rng(1234).integers(0, 2, (B, 256)) * 2 - 1).astype(np.int8)).cuda()
run = (lambda: be.evaluate_grad(pats)) if len(sys.argv) > 2 and sys.argv[2] == "grad" else (lambda: be.evaluate(pats))
run(); torch.cuda.synchronize(); be.debug_fetch("phase_clk", 1)
run(); torch.cuda.synchronize()
c = be.debug_fetch("phase_clk", 1)
tot = c.sum()
for i, v in enumerate(c):
    if v: print(f"phase {i:2d}: {v / B:12.0f} clk/structure  {100 * v / tot:5.1f}%")
print(f"total {tot / B:.0f} clk per structure (per CTA)")
