import sys, numpy as np
sys.path.insert(0, ".")
from physics_informed_reinforcement_learning_b200 import RCWABatch
from oracle import rcwa_oracle as O
for n_cells, order in ((20000, 40), (65536, 40), (12000, 60), (3000, 100)):
    rng = np.random.default_rng(n_cells)
    pat = np.repeat(rng.choice([-1.0, 1.0], -(-n_cells // 50)), 50)[:n_cells]
    be = RCWABatch(n_cells, order, max_batch=2)
    eff, g, st = be.evaluate_grad_host(pat[None, :], 1100.0, 60.0, 325.0)
    idx = [0, n_cells // 3, n_cells - 1]
    fd = O.fd_gradient(pat, idx, order=order)
    print(n_cells, order, "eff", eff[0], "status", st, "grad", g[0, idx], "fd", fd, "err", np.abs(g[0, idx] - fd).max(), flush=True)
    be.close()
