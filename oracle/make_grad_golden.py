"""Golden vectors for the adjoint gradient (test infrastructure only).

The reference has no gradient code: the oracle is the central finite difference of the pinned forward
restatement (oracle/rcwa_oracle.py::fd_gradient, h = 1e-4) w.r.t. one continuous cell value.  This script writes
tests/golden/grad_golden.npz:  for each case the pattern, the parameters, the efficiency and the full FD gradient.

    python -m oracle.make_grad_golden          (about 2 minutes on 8 cores)
"""
import os
from concurrent.futures import ProcessPoolExecutor

import numpy as np

from oracle import rcwa_oracle as O

CASES = {
    # name: (pattern, wavelength, angle, order, thickness)
    "G7_blocks16_N40": (np.repeat(np.random.default_rng(3).choice([-1., 1.], 16), 16), 1100.0, 60.0, 40, 325.0),
    "G1_rand256_N40": (np.random.default_rng(0).choice([-1., 1.], 256), 1100.0, 60.0, 40, 325.0),
    "G4_half_N40": (np.array([1.] * 128 + [-1.] * 128), 1100.0, 60.0, 40, 325.0),
    "G7_blocks16_1000nm_65deg_N60": (np.repeat(np.random.default_rng(3).choice([-1., 1.], 16), 16), 1000.0, 65.0, 60, 325.0),
    "blocks8_900nm_50deg_N40": (np.repeat(np.random.default_rng(8).choice([-1., 1.], 32), 8), 900.0, 50.0, 40, 325.0),
}


def _one(args):
    pat, i, wl, ang, order, thick = args
    return O.fd_gradient(pat, [i], wl, ang, order, thick)[0]


def main():
    os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
    out = {}
    with ProcessPoolExecutor(os.cpu_count()) as ex:
        for name, (pat, wl, ang, order, thick) in CASES.items():
            g = np.array(list(ex.map(_one, [(pat, i, wl, ang, order, thick) for i in range(pat.size)], chunksize=8)))
            out[name + "/pattern"] = pat
            out[name + "/params"] = np.array([wl, ang, order, thick])
            out[name + "/eff"] = O.efficiency(pat, wl, ang, order, thick)
            out[name + "/fd_grad"] = g
            print(name, out[name + "/eff"], np.abs(g).max())
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "grad_golden.npz"), **out)


if __name__ == "__main__":
    main()
