"""Debug helper: gradient call at order 45 (ff = 91: shared-memory g_tail with one spilled matrix)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from physics_informed_reinforcement_learning_b200 import RCWABatch
order = int(sys.argv[1]) if len(sys.argv) > 1 else 45
pat = np.random.default_rng(order).choice([-1.0, 1.0], 128)
be = RCWABatch(128, order, max_batch=4, device=0)
print("forward", be.evaluate_host(pat[None, :].astype(np.int8), 1000.0, 55.0, 300.0))
eff, grad, st = be.evaluate_grad_host(pat[None, :], 1000.0, 55.0, 300.0)
print(eff, st, grad[0, :4])
