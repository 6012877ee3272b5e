import sys; sys.path.insert(0,'.')
import numpy as np, torch
from physics_informed_reinforcement_learning_b200 import RCWABatch
for order in (40, 60):
    be = RCWABatch(256, order, max_batch=2048); be.set_subchunk(0)
    pats = torch.from_numpy((np.random.default_rng(1234).integers(0, 2, (2048, 256)) * 2 - 1).astype(np.int8)).cuda()
    be.evaluate(pats); torch.cuda.synchronize()
    ws = be.debug_fetch("wstatus", 2048).astype(int)
    lam = be.debug_fetch("lam", 2048)
    fb = np.nonzero(ws & 0x100)[0]
    print(order, "fallbacks", len(fb), fb[:10])
    k0d = 2*np.pi/1100*325
    for b in fb[:5]:
        l = lam[b]; pos = l[l > 0]; X = np.exp(-k0d*np.sqrt(pos))
        print("  np", (l <= 0).sum(), "ns", (X >= 1e-11).sum(), "ni", (X < 1e-11).sum())
