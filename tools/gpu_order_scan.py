"""Forward parity scan over orders: python tools/gpu_order_scan.py lo hi  (prints eff error vs the oracle and the status)"""
import sys
import numpy as np
sys.path.insert(0, ".")
from physics_informed_reinforcement_learning_b200 import RCWABatch
from oracle import rcwa_oracle as O
lo, hi = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(4242)
n_cells, wl, ang, thick = 96, 1050.0, 58.0, 310.0
for order in range(lo, hi + 1):
    pats = np.stack([rng.choice([-1, 1], n_cells), np.repeat(rng.choice([-1, 1], n_cells // 8), 8)]).astype(np.int8)
    be = RCWABatch(n_cells, order, max_batch=4)
    eff, r3, t3, st = be.evaluate_host(pats, wl, ang, thick, return_orders=True)
    ref = [O.evaluate(p.astype(np.float64), wl, ang, order, thick).eff for p in pats]
    ws = be.debug_fetch("wstatus", 2) if hasattr(be, "debug_fetch") else None
    print(order, 2 * order + 1, "err", np.abs(eff - np.array(ref)), "status", st, "wstatus", ws, flush=True)
    be.close()
