"""Cluster-path check on the GPU box: parity vs the oracle at a few large orders, per-kernel times, phase clocks.
usage: python tools/gpu_cluster_check.py [batch_for_timing]"""
import os, sys, time
os.environ["PIRL_PHASE_CLOCKS"] = "1"
sys.path.insert(0, ".")
import numpy as np, torch
from physics_informed_reinforcement_learning_b200 import RCWABatch
from oracle import rcwa_oracle as O

rng = np.random.default_rng(77)


def parity(n_cells, order, nb=6, wl=1100.0, ang=60.0, thick=325.0):
    pats = rng.choice([-1, 1], (nb, n_cells)).astype(np.int8)
    pats[1] = np.repeat(rng.choice([-1, 1], n_cells // 16), 16)
    pats[2] = np.repeat(rng.choice([-1, 1], n_cells // 32), 32)
    be = RCWABatch(n_cells, order, max_batch=max(nb, 8))
    t0 = time.time()
    eff, r3, t3, st = be.evaluate_host(pats, wl, ang, thick, return_orders=True)
    ref = np.array([O.evaluate(p.astype(np.float64), wl, ang, order, thick).eff for p in pats])
    print(f"n={n_cells} order={order} ff={2*order+1}: max|err|={np.abs(eff-ref).max():.2e} eff={eff[:3]} status={st} ({time.time()-t0:.1f}s)", flush=True)
    be.close()
    return np.abs(eff - ref).max()


if os.environ.get("PIRL_FORCE_CLUSTER"):
    parity(256, 40)
    parity(96, 17)
    sys.exit(0)

worst = 0.0
for n_cells, order in ((256, 41), (256, 64), (256, 75), (512, 100), (512, 103), (256, 90)):
    worst = max(worst, parity(n_cells, order))
print("worst", worst)

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
be = RCWABatch(512, 100, max_batch=B)
pats = torch.from_numpy((np.random.default_rng(5).integers(0, 2, (B, 512)) * 2 - 1).astype(np.int8)).cuda()
be.evaluate(pats); torch.cuda.synchronize(); be.debug_fetch("phase_clk", 1)
be.set_stage_timing(True); be.stage_times()
t0 = time.time()
be.evaluate(pats); torch.cuda.synchronize()
dt = time.time() - t0
print(f"B={B} order 100: {dt*1e3:.1f} ms wall -> {B/dt:.0f} evals/s")
print(be.stage_times())
c = be.debug_fetch("phase_clk", 1)
for i, v in enumerate(c):
    if v: print(f"phase {i:2d}: {v / B:12.0f} clk/structure")
