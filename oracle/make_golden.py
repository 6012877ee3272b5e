"""Generate tests/golden/*.npz from the UNMODIFIED reference solver (authoring container only).

    python oracle/make_golden.py

Writes
  physics_informed_reinforcement_learning_b200/data/p_si_nk.npz   material table (data) from nk_data/matlab/p_Si.mat
  tests/golden/rcwa_golden.npz     named cases G1..G8 + Reticolo known answer (reference outputs)
  tests/golden/rcwa_batch.npz      256 mixed random/blocky patterns, N=40, 1100 nm / 60 deg (reference eff, cuts)
  tests/golden/rcwa_sweep.npz      small (wavelength, angle) grid, N=60 (config C4 sample) + FD gradients (G9)
  tests/golden/env_golden.npz      random_bunch_init draws under np.random.seed(...) (function source exec'd from the reference)

Every value stored under a key starting with ``ref_`` comes from the reference code itself
(``oracle/ref_bridge.py``), not from ``oracle/rcwa_oracle.py``.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_bridge  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
DATA = os.path.join(ROOT, "physics_informed_reinforcement_learning_b200", "data")


def mixed_patterns(rng, count, n_cells):
    """Noise + blocky patterns so efficiencies span 0..0.9."""
    pats = []
    for i in range(count):
        kind = i % 4
        if kind == 0:
            p = rng.choice([-1.0, 1.0], n_cells)
        else:
            blk = (8, 16, 32)[kind - 1]
            p = np.repeat(rng.choice([-1.0, 1.0], n_cells // blk), blk)
        pats.append(p)
    return np.array(pats)


def main():
    os.makedirs(GOLD, exist_ok=True)
    os.makedirs(DATA, exist_ok=True)
    assert ref_bridge.available(), "needs /root/reference"

    wl, n, k = ref_bridge.reference_material_table()
    np.savez(os.path.join(DATA, "p_si_nk.npz"), wl=wl, n=n, k=k)

    # ---- named cases -------------------------------------------------------------------
    cases = {}

    def add(name, pattern, wavelength=1100, angle=60, order=40, thickness=325, algo="TMM"):
        t = time.time()
        eff, rc, tc = ref_bridge.reference_evaluate(pattern, wavelength, angle, order, thickness, algo)
        cases[name] = dict(pattern=np.asarray(pattern, dtype=np.float64), wavelength=wavelength, angle=angle,
                           order=order, thickness=thickness, ref_eff=eff, ref_refl_cut=rc, ref_tran_cut=tc)
        print(f"{name:24s} eff={eff!r}  ({time.time() - t:.2f}s)")

    g1 = np.random.default_rng(0).choice([-1., 1.], 256)
    add("G1_rand256_N40", g1)
    add("G1_rand256_N40_SMM", g1, algo="SMM")
    add("G2_rand256_N60", g1, order=60)
    add("G3_all_plus", np.ones(256))
    add("G3_all_minus", -np.ones(256))
    half = np.array([1.] * 128 + [-1.] * 128)
    add("G4_half_N40", half)
    add("G4_half_N60", half, order=60)
    np.random.seed(42)
    ns = {}
    with open(os.path.join(ref_bridge.REF_ROOT, "deflector-gym", "deflector_gym", "envs", "utils.py")) as fh:
        src = fh.read()
    start = src.index("def random_bunch_init")
    exec("import numpy as np\n" + src[start:], ns)  # the reference's own function object
    ref_bunch = ns["random_bunch_init"]
    g5 = ref_bunch()
    add("G5_bunch_seed42", g5)
    add("G6_rand512_N100", np.random.default_rng(7).choice([-1., 1.], 512), order=100)
    g7 = np.repeat(np.random.default_rng(3).choice([-1., 1.], 16), 16)
    add("G7_block16_N40", g7)
    add("G7_block16_1000nm_65deg_N60", g7, wavelength=1000, angle=65, order=60)
    g8 = g7.copy(); g8[100] = -g8[100]
    add("G8_block16_flip100", g8)
    ka = ref_bridge.reticolo_known_answer()
    add("KA_reticolo_800nm_70deg", ka["pattern"], wavelength=int(ka["wavelength"]), angle=int(ka["angle"]),
        thickness=int(ka["thickness"]))
    cases["KA_reticolo_800nm_70deg"]["reticolo_efficiency"] = ka["efficiency"]
    cases["KA_reticolo_800nm_70deg"]["reticolo_n_Si"] = ka["n_Si"]
    add("KA_reversed", ka["pattern"][::-1].copy(), wavelength=800, angle=70)

    flat = {}
    for name, d in cases.items():
        for kk, v in d.items():
            flat[f"{name}/{kk}"] = np.asarray(v)
    np.savez_compressed(os.path.join(GOLD, "rcwa_golden.npz"), **flat)

    # ---- batch of 256 mixed patterns, headline geometry ------------------------------------
    rng = np.random.default_rng(20261104)
    pats = mixed_patterns(rng, 256, 256)
    effs, rcs, tcs = [], [], []
    t = time.time()
    for p in pats:
        e, rc, tc = ref_bridge.reference_evaluate(p, 1100, 60, 40, 325)
        effs.append(e); rcs.append(rc); tcs.append(tc)
    print(f"batch of {len(pats)}: {time.time() - t:.1f}s, eff range {min(effs):.4g}..{max(effs):.4g}")
    np.savez_compressed(os.path.join(GOLD, "rcwa_batch.npz"), patterns=pats.astype(np.int8),
                        wavelength=1100, angle=60, order=40, thickness=325,
                        ref_eff=np.array(effs), ref_refl_cut=np.array(rcs), ref_tran_cut=np.array(tcs))

    # ---- sweep sample (C4) + FD gradient (G9) --------------------------------------------
    rng = np.random.default_rng(4)
    spats = mixed_patterns(rng, 4, 256)
    wls = [900, 1000, 1100, 1200]
    angs = [50, 60, 75]
    grid = []
    for w in wls:
        for a in angs:
            for ip, p in enumerate(spats):
                e, rc, tc = ref_bridge.reference_evaluate(p, w, a, 60, 325)
                grid.append((w, a, ip, e))
    grid = np.array(grid)
    idx = np.array([5, 100, 200])
    h = 1e-4
    fd = []
    for i in idx:
        sp = g7.copy(); sp[i] += h
        sm = g7.copy(); sm[i] -= h
        fd.append((ref_bridge.reference_evaluate(sp)[0] - ref_bridge.reference_evaluate(sm)[0]) / (2 * h))
    print("FD grads on G7:", fd)
    np.savez_compressed(os.path.join(GOLD, "rcwa_sweep.npz"), patterns=spats.astype(np.int8), order=60,
                        thickness=325, grid=grid, fd_pattern=g7, fd_idx=idx, ref_fd_grad=np.array(fd), fd_h=h)

    # ---- env fixtures ------------------------------------------------------------------
    seeds = np.arange(20)
    inits = []
    for s in seeds:
        np.random.seed(int(s))
        inits.append(ref_bunch())
    np.savez_compressed(os.path.join(GOLD, "env_golden.npz"), seeds=seeds, ref_bunch=np.array(inits).astype(np.int8))
    print("done")


if __name__ == "__main__":
    main()
