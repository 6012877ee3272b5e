"""CTA timeline of one evaluate() call (PIRL_TRACE=1): per kernel start/end, SMs used, co-residency.  Run on the GPU box."""
import os, sys
os.environ["PIRL_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from physics_informed_reinforcement_learning_b200 import RCWABatch
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
be = RCWABatch(256, 40, max_batch=B)
pats = torch.from_numpy((np.random.default_rng(1234).integers(0, 2, (B, 256)) * 2 - 1).astype(np.int8)).cuda()
for _ in range(2):
    be.evaluate(pats); torch.cuda.synchronize(); be.debug_fetch("trace", 1)
be.evaluate(pats); torch.cuda.synchronize()
tr = be.debug_fetch("trace", 1)
kid = (tr[:, 0] >> np.uint64(32)).astype(int); sm = (tr[:, 0] & np.uint64(0xffffffff)).astype(int)
t0 = tr[:, 2].astype(np.int64); t1 = tr[:, 3].astype(np.int64)
base = t0.min()
names = {2: "k_ql", 3: "k_apply", 5: "rk_front", 6: "rk_tail"}
print(f"{len(tr)} CTA records, span {(t1.max() - base) / 1e6:.3f} ms")
# group launches: records of the same kernel separated in time -> cluster by start gaps
for k in sorted(set(kid)):
    idx = np.where(kid == k)[0]
    order = idx[np.argsort(t0[idx])]
    print(f"== {names.get(k, k)}: {len(idx)} CTAs, first start {(t0[idx].min() - base) / 1e6:.3f} ms, last end {(t1[idx].max() - base) / 1e6:.3f} ms, "
          f"mean CTA duration {(t1[idx] - t0[idx]).mean() / 1e3:.1f} us, SMs used {len(set(sm[idx]))}")
# timeline in 0.25 ms bins: number of CTAs of each kernel alive
edges = np.arange(0, (t1.max() - base) / 1e6 + 0.25, 0.25)
print("t(ms)   " + "  ".join(f"{names.get(k, k):>9s}" for k in sorted(set(kid))))
for e in edges:
    t = base + int(e * 1e6)
    print(f"{e:6.2f}  " + "  ".join(f"{int(((t0 <= t) & (t1 > t) & (kid == k)).sum()):9d}" for k in sorted(set(kid))))
