"""Forward (and gradient) parity scan over cell counts and layer thicknesses at a fixed order."""
import sys
import numpy as np
sys.path.insert(0, ".")
from physics_informed_reinforcement_learning_b200 import RCWABatch
from oracle import rcwa_oracle as O
order = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(99)
for n_cells in ((2, 3, 7, 64, 100, 255, 257, 1000, 4096, 20000, 65536) if len(sys.argv) < 3 else (9000, 12000, 20000, 65536)):
    pats = np.stack([rng.choice([-1, 1], n_cells), np.repeat(rng.choice([-1, 1], -(-n_cells // 5)), 5)[:n_cells]]).astype(np.int8)
    try:
        be = RCWABatch(n_cells, order, max_batch=4)
        eff, r3, t3, st = be.evaluate_host(pats, 1100.0, 60.0, 325.0, return_orders=True)
        ref = np.array([O.evaluate(p.astype(np.float64), 1100.0, 60.0, order, 325.0).eff for p in pats])
        msg = f"err {np.abs(eff - ref).max():.2e} status {st}"
        if n_cells <= 4096:
            e2, g, st2 = be.evaluate_grad_host(pats.astype(np.float64), 1100.0, 60.0, 325.0)
            idx = [0, n_cells // 2, n_cells - 1]
            fd = O.fd_gradient(pats[1].astype(np.float64), idx, order=order)
            msg += f"  grad err {np.abs(g[1, idx] - fd).max():.2e} (|fd| {np.abs(fd).max():.2e}) eff diff {np.abs(e2 - eff).max():.1e}"
        be.close()
    except Exception as ex:  # noqa: BLE001
        msg = f"EXC {type(ex).__name__}: {str(ex)[:150]}"
    print("n_cells", n_cells, msg, flush=True)
if len(sys.argv) >= 3: sys.exit(0)
pats = np.repeat(rng.choice([-1, 1], (1, 32)), 8, axis=1).astype(np.int8)
be = RCWABatch(256, order, max_batch=4)
for thick in (1.0, 10.0, 100.0, 1000.0, 5000.0, 50000.0):
    for wl, ang in ((1100.0, 60.0), (400.0, 20.0), (2000.0, 85.0), (1100.0, 5.0)):
        eff, r3, t3, st = be.evaluate_host(pats, wl, ang, thick, return_orders=True)
        ref = O.evaluate(pats[0].astype(np.float64), wl, ang, order, thick).eff
        print(f"thick {thick} wl {wl} ang {ang}: eff {eff[0]:.6e} ref {ref:.6e} err {abs(eff[0] - ref):.2e} status {st}", flush=True)
