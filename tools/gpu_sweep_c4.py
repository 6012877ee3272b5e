"""Config C4 of SURVEY.md section 8(d) on one GPU: wavelength 900..1200 step 50 (7) x angle 50..75 step 5 (6) x 256 patterns
of default_rng(4), 256 cells, order 60, forward and forward + gradient, through sweep.py (one batch axis of 10752 structures
with per-structure parameter rows).  Times each with CUDA events (3 warm-ups, 5 timed) and checks a few cells against the
oracle.   usage:  python tools/gpu_sweep_c4.py [max_batch]"""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from physics_informed_reinforcement_learning_b200 import RCWABatch
from physics_informed_reinforcement_learning_b200.sweep import sweep
from oracle import rcwa_oracle as O

max_batch = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
wls = np.arange(900.0, 1201.0, 50.0)
angs = np.arange(50.0, 76.0, 5.0)
rng = np.random.default_rng(4)
pats = rng.choice([-1, 1], (256, 256)).astype(np.int8)
pats[::4] = np.repeat(rng.choice([-1, 1], (64, 16)), 16, axis=1)          # blocky rows: efficiencies beyond the noise floor
dpat = torch.from_numpy(pats).cuda()
be = RCWABatch(256, 60, max_batch=max_batch, device=0)
cells = pats.shape[0] * wls.size * angs.size


def timed(fn, warm=3, reps=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); out = fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return out, float(np.median(ts)), float(min(ts))


eff, med_f, best_f = timed(lambda: sweep(be, dpat, wls, angs))
(effg, grad, st), med_g, best_g = timed(lambda: sweep(be, dpat, wls, angs, grad=True))
assert int(st.abs().sum()) == 0 and float((eff - effg).abs().max()) <= 1e-10
e = eff.cpu().numpy(); g = grad.cpu().numpy()
worst = worst_g = 0.0
t0 = time.time()
for (p, i, j) in [(0, 0, 0), (4, 3, 2), (17, 6, 5), (128, 2, 4), (255, 5, 1), (64, 4, 3)]:
    ref = O.efficiency(pats[p].astype(np.float64), wls[i], angs[j], 60, 325.0)
    worst = max(worst, abs(e[p, i, j] - ref))
    idx = [7, 130, 249]
    fd = O.fd_gradient(pats[p].astype(np.float64), idx, wls[i], angs[j], 60, 325.0)
    worst_g = max(worst_g, float(np.abs(g[p, i, j, idx] - fd).max() / (1e-8 + 1e-5 * np.abs(fd)).min()))
print(f"C4: {cells} structures (256 patterns x {wls.size} wavelengths x {angs.size} angles), order 60, max_batch {max_batch}")
print(f"forward          : median {med_f:.1f} ms, best {best_f:.1f} ms -> {cells / med_f * 1e3:.0f} evals/s")
print(f"forward+gradient : median {med_g:.1f} ms, best {best_g:.1f} ms -> {cells / med_g * 1e3:.0f} gradients/s")
print(f"efficiency range {e.min():.3e} .. {e.max():.4f}; max |eff - oracle| over 6 cells {worst:.2e}; "
      f"gradient error / tolerance (1e-8 + 1e-5 |fd|) max {worst_g:.3f}; oracle time {time.time() - t0:.1f} s")
assert worst <= 1e-6 and worst_g <= 1.0
be.close()
