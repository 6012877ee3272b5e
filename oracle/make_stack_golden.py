"""Write tests/golden/rcwa_stack_golden.npz from the UNMODIFIED reference (authoring container only): multilayer stacks and
TE polarisation, 'TMM' and 'SMM', through JLABCode.reproduce_acs_cell (JLAB/solver.py:16-38) with ucell of shape (L, 1, n).
    python -m oracle.make_stack_golden
"""
import os
import numpy as np
from oracle import ref_bridge

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "rcwa_stack_golden.npz")


def ref(structs, wl, ang, order, th, pol, algo):
    J = ref_bridge._load()
    wls = np.array([float(wl)])
    period = abs(wls / np.sin(ang / 180 * np.pi))
    c = J(grating_type=0, n_I=1.45, n_II=1., theta=0, phi=0., fourier_order=order, period=period, wls=wls, pol=pol,
          patterns=None, ucell=np.asarray(structs, float)[:, None, :].copy(), thickness=np.asarray(th, float), algo=algo)
    eff, refl, tran = c.reproduce_acs_cell('p_si__real', 1)
    return float(eff), np.asarray(refl, float), np.asarray(tran, float)


def main():
    rng = np.random.default_rng(2024)
    cases = {}
    def add(name, structs, wl, ang, order, th, pol, algo):
        eff, refl, tran = ref(structs, wl, ang, order, th, pol, algo)
        for k, v in dict(structs=np.asarray(structs, float), wavelength=wl, angle=ang, order=order, thickness=np.asarray(th, float),
                         pol=pol, algo=algo, ref_eff=eff, ref_refl_cut=refl, ref_tran_cut=tran).items():
            cases[f"{name}/{k}"] = np.asarray(v)
        print(f"{name:28s} eff {eff:.15f}")
    blocks2 = np.repeat(rng.choice([-1., 1.], (2, 16)), 16, axis=1)
    blocks3 = np.repeat(rng.choice([-1., 1.], (3, 32)), 8, axis=1)
    rnd1 = rng.choice([-1., 1.], (1, 256))
    for algo in ("TMM", "SMM"):
        add(f"ML2_block16_N20_{algo}", blocks2, 1100., 60., 20, [200., 325.], 1, algo)
        add(f"ML3_block8_N30_{algo}", blocks3, 1000., 65., 30, [150., 90., 260.], 1, algo)
        add(f"TE1_rand256_N40_{algo}", rnd1, 1100., 60., 40, [325.], 0, algo)
        add(f"TE1_block16_N20_{algo}", blocks2[:1], 900., 50., 20, [325.], 0, algo)
        add(f"TE2_block16_N20_{algo}", blocks2, 1100., 60., 20, [200., 325.], 0, algo)
    np.savez_compressed(OUT, **cases)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
