"""GPU check of pirl_rcwa_eval_grad against the NumPy design prototype (tools/proto_grad.py), full gradient vectors."""
import sys, time
import numpy as np
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tools")
from physics_informed_reinforcement_learning_b200 import RCWABatch
from oracle import rcwa_oracle as O
import proto_grad as PG

def run(n_cells, order, pats, wl=1100.0, ang=60.0):
    be = RCWABatch(n_cells, order, max_batch=max(64, len(pats)), device=0)
    eff, grad, st = be.evaluate_grad(torch.from_numpy(pats).cuda(), wl, ang, 325.0)
    torch.cuda.synchronize()
    eff = eff.cpu().numpy(); grad = grad.cpu().numpy()
    effh, gradh, sth = be.evaluate_grad_host(pats, wl, ang, 325.0)
    print("host==device", np.array_equal(effh, eff), np.array_equal(gradh, grad), "status", st.cpu().numpy().tolist())
    nr = O.n_silicon_real(wl)
    for i, p in enumerate(pats):
        e, g = PG.forward_backward(p.astype(float), wl, ang, order, 325.0, n_ridge=nr)
        print(f"N={order} pat{i}: eff {eff[i]:.12f} proto {e:.12f}  max|g| {np.abs(g).max():.3e}  max|dg| {np.abs(g - grad[i]).max():.3e}")
    return be

rng = np.random.default_rng(3)
pats = np.stack([np.repeat(rng.choice([-1, 1], 16), 16), np.random.default_rng(0).choice([-1, 1], 256),
                 np.array([1] * 128 + [-1] * 128), np.ones(256, int), np.repeat(np.random.default_rng(8).choice([-1, 1], 32), 8)]).astype(np.int8)
be = run(256, 40, pats)
run(256, 60, pats[:3], 1000.0, 65.0)
if len(sys.argv) > 1:
    p512 = np.random.default_rng(7).choice([-1, 1], (2, 512)).astype(np.int8)
    run(512, 100, p512)
# timing at batch 4096
big = torch.from_numpy(np.random.default_rng(1234).integers(0, 2, (4096, 256)).astype(np.int8) * 2 - 1).cuda()
be = RCWABatch(256, 40, max_batch=4096, device=0)
for _ in range(2): be.evaluate_grad(big)
torch.cuda.synchronize()
be.set_stage_timing(True)
t0 = time.time()
for _ in range(3): be.evaluate_grad(big)
torch.cuda.synchronize()
dt = (time.time() - t0) / 3
print(f"eval_grad 4096: {dt * 1e3:.2f} ms -> {4096 / dt:.0f} grads/s", {k: round(v[0] / 3, 3) for k, v in be.stage_times().items() if v[1]})
