"""Small forward + gradient calls across the kernel variants (register-tile, padded, shared-memory, generic front):
statuses zero, gradients finite, forward and gradient entry points agree on the efficiency.  (compute-sanitizer is closed
on this pool, so this runs plain.)   python tools/gpu_variant_smoke.py"""
import sys
import numpy as np
sys.path.insert(0, ".")
from physics_informed_reinforcement_learning_b200 import RCWABatch
for n_cells, order, B in ((256, 40, 5), (256, 60, 3), (256, 45, 3), (96, 60, 2), (256, 8, 3), (512, 100, 2), (5000, 40, 2)):
    rng = np.random.default_rng(order)
    pats = rng.choice([-1, 1], (B, n_cells)).astype(np.int8)
    be = RCWABatch(n_cells, order, max_batch=4)
    eff = be.evaluate_host(pats)
    eff_g, g, st = be.evaluate_grad_host(pats.astype(np.float64))
    print(n_cells, order, "eff", eff[:2], "grad-path eff diff", np.abs(eff - eff_g).max(), "status", st, flush=True)
    assert np.isfinite(g).all() and not st.any()
    be.close()
print("variant smoke done")
