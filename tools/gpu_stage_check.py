"""GPU bring-up check: compares every pipeline stage of the CUDA path with tools/proto_kernels.py (NumPy)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import proto_kernels as PK
from physics_informed_reinforcement_learning_b200 import RCWABatch, n_silicon_real

def run(order, n_cells, B=6, wl=1100., ang=60., seed=0):
    rng = np.random.default_rng(seed)
    pats = rng.choice([-1, 1], (B, n_cells)).astype(np.int8)
    pats[1] = np.repeat(rng.choice([-1, 1], n_cells // 16), 16)
    pats[2] = 1
    if B > 3:
        pats[3] = -1
    be = RCWABatch(n_cells, order, max_batch=B)
    t = torch.from_numpy(pats).cuda()
    eff, r3, t3, st = be.evaluate(t, wl, ang, 325., return_orders=True)
    torch.cuda.synchronize()
    eff = eff.cpu().numpy(); r3 = r3.cpu().numpy(); t3 = t3.cpu().numpy(); st = st.cpu().numpy()
    Cg = be.debug_fetch("C", B); tri = be.debug_fetch("tri", B); lam = be.debug_fetch("lam", B)
    W = be.debug_fetch("W", B); nrot = be.debug_fetch("nrot", B)
    nr = float(n_silicon_real(wl))
    print(f"== order {order} n_cells {n_cells}  status {st}  nrot {nrot}")
    for b in range(B):
        e, der, det, S = PK.evaluate(pats[b].astype(float), wl, ang, order, 325., n_ridge=nr, stages=True)
        dC = np.abs(Cg[b] - S['C']).max()
        dd = np.abs(tri[b, 0] - S['d']).max(); de = np.abs(np.abs(tri[b, 1][:-1]) - np.abs(S['e'])).max()
        dl = np.abs(np.sort(lam[b]) - np.sort(S['lam'])).max()
        orth = np.abs(W[b].T @ S['A'] @ W[b] - np.eye(2 * order + 1)).max()
        print(f"  b={b} dC={dC:.1e} dd={dd:.1e} de={de:.1e} dlam={dl:.1e} orth={orth:.1e} "
              f"eff gpu={eff[b]:.15g} ref={e:.15g} d={abs(eff[b]-e):.1e} dr3={np.abs(r3[b]-der).max():.1e} dt3={np.abs(t3[b]-det).max():.1e}")
    be.close()

if __name__ == '__main__':
    run(40, 256)
    run(60, 256, B=4)
    run(100, 512, B=3)
    run(8, 64, B=4)
