import sys
import numpy as np, torch
sys.path.insert(0, "."); sys.path.insert(0, "tools")
from physics_informed_reinforcement_learning_b200 import RCWABatch
import proto_grad as PG
rng = np.random.default_rng(1234)
pats = (rng.integers(0, 2, (4096, 256)) * 2 - 1).astype(np.int8)
pats[:256] = np.repeat(rng.integers(0, 2, (256, 16)) * 2 - 1, 16, axis=1)
be = RCWABatch(256, 40, max_batch=4096, device=0)
d = torch.from_numpy(pats).cuda()
eff, grad, st = be.evaluate_grad(d)
eff_r, grad_r, _ = be.evaluate_grad(torch.roll(d, 37, dims=1))
err = (grad_r - torch.roll(grad, 37, dims=1)).abs().amax(dim=1).cpu().numpy()
order = np.argsort(-err)[:6]
print("worst structures", order, err[order])
g = grad.cpu().numpy(); gr = grad_r.cpu().numpy()
for b in order[:3]:
    e, gp = PG.forward_backward(pats[b].astype(float))
    e2, gp2 = PG.forward_backward(np.roll(pats[b], 37).astype(float))
    print(b, "eff", float(eff[b]), e, "| gpu-proto", np.abs(g[b] - gp).max(), "| gpu_rolled-proto_rolled", np.abs(gr[b] - gp2).max(),
          "| proto equivariance", np.abs(np.roll(gp, 37) - gp2).max(), "max|g|", np.abs(gp).max())
    lam = np.sort(be.debug_fetch("lam", 4096)[b]) if False else None
