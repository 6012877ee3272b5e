"""CPU: the oracle (oracle/rcwa_oracle.py) against fixtures produced by the UNMODIFIED reference solver
(oracle/make_golden.py) and against the recorded Reticolo known answer."""
import os

import numpy as np
import pytest

from oracle import rcwa_oracle as O


def _cases(golden_dir):
    g = np.load(os.path.join(golden_dir, "rcwa_golden.npz"))
    names = sorted({k.split("/")[0] for k in g.files})
    return g, names


def test_named_cases_match_reference(golden_dir):
    g, names = _cases(golden_dir)
    assert len(names) >= 14
    for n in names:
        if n.endswith("_SMM"):
            continue  # same pattern solved by the reference's S-matrix variant; checked below
        r = O.evaluate(g[n + "/pattern"], float(g[n + "/wavelength"]), float(g[n + "/angle"]), int(g[n + "/order"]),
                       float(g[n + "/thickness"]))
        assert abs(r.eff - float(g[n + "/ref_eff"])) <= 1e-12, n
        np.testing.assert_allclose(r.tran_cut, g[n + "/ref_tran_cut"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(r.refl_cut, g[n + "/ref_refl_cut"], rtol=0, atol=1e-12)


def test_smm_and_tmm_reference_agree(golden_dir):
    g, _ = _cases(golden_dir)
    assert abs(float(g["G1_rand256_N40/ref_eff"]) - float(g["G1_rand256_N40_SMM/ref_eff"])) < 1e-12


def test_survey_pinned_values(golden_dir):
    """Values quoted in SURVEY.md section 8c (independent run of the reference)."""
    g, _ = _cases(golden_dir)
    assert float(g["G1_rand256_N40/ref_eff"]) == pytest.approx(0.0004665720258770959, abs=1e-15)
    assert float(g["G7_block16_N40/ref_eff"]) == pytest.approx(0.055298772147661227, abs=1e-15)
    assert float(g["G4_half_N40/ref_eff"]) == pytest.approx(0.2498360217616929, abs=1e-14)
    assert float(g["G6_rand512_N100/ref_eff"]) == pytest.approx(0.0006183899938941343, abs=1e-14)


def test_reticolo_known_answer(golden_dir):
    g, _ = _cases(golden_dir)
    n = "KA_reticolo_800nm_70deg"
    r = O.evaluate(g[n + "/pattern"], 800.0, 70.0, 40, 325.0)
    assert abs(r.eff - float(g[n + "/reticolo_efficiency"])) < 5e-8      # float32 rounding of the recorded value
    assert abs(O.n_silicon_real(800.0) - float(g[n + "/reticolo_n_Si"])) < 1e-6
    rev = O.evaluate(g[n + "/pattern"][::-1], 800.0, 70.0, 40, 325.0)
    assert abs(rev.eff - 0.027283368919125597) < 1e-12                   # orientation pin


def test_energy_conservation_and_trivial_patterns():
    rng = np.random.default_rng(5)
    p = np.repeat(rng.choice([-1.0, 1.0], 32), 8)
    r = O.evaluate(p)
    assert abs(r.de_ri.sum() + r.de_ti.sum() - 1.0) < 1e-10   # lossless
    assert O.efficiency(np.ones(256)) == pytest.approx(0.0, abs=1e-20)
    assert O.efficiency(-np.ones(256)) == pytest.approx(0.0, abs=1e-20)


def test_batch_fixture_sample(golden_dir):
    b = np.load(os.path.join(golden_dir, "rcwa_batch.npz"))
    idx = [0, 1, 2, 3, 77, 255]
    got = O.efficiency_batch(b["patterns"][idx].astype(np.float64))
    np.testing.assert_allclose(got, b["ref_eff"][idx], rtol=0, atol=1e-12)
    assert b["ref_eff"].max() > 0.5   # fixture spans the efficiency range, not just noise-level values


def test_fd_gradient_fixture(golden_dir):
    s = np.load(os.path.join(golden_dir, "rcwa_sweep.npz"))
    got = O.fd_gradient(s["fd_pattern"], s["fd_idx"], h=float(s["fd_h"]))
    np.testing.assert_allclose(got, s["ref_fd_grad"], rtol=1e-6, atol=1e-10)


def test_material_interpolation():
    assert O.n_silicon_real(1100.0) == pytest.approx(3.551726470588235, abs=1e-14)


def test_grad_golden_consistent_with_oracle_fd(golden_dir):
    """tests/golden/grad_golden.npz (oracle/make_grad_golden.py) is the oracle's own finite difference; spot-check it and
    the values recorded in SURVEY.md section 8c (G9)."""
    g = np.load(os.path.join(golden_dir, "grad_golden.npz"))
    n = "G7_blocks16_N40"
    idx = [5, 100, 200]
    got = O.fd_gradient(g[n + "/pattern"], idx)
    np.testing.assert_allclose(got, g[n + "/fd_grad"][idx], rtol=2e-6, atol=2e-9)    # FD rounding noise ~ eps/h
    np.testing.assert_allclose(got, [-7.3443880e-04, 3.6901669e-03, -1.6564464e-03], rtol=2e-6)
    assert abs(O.efficiency(g[n + "/pattern"]) - float(g[n + "/eff"])) < 1e-12       # BLAS thread count changes the last digits


def test_stack_restatement_matches_reference_fixtures(golden_dir):
    """oracle.rcwa_oracle.solve_stack (multilayer, TE, 'TMM' and 'SMM') against the unmodified reference (fixtures written by
    oracle/make_stack_golden.py); the reference's own SMM / TMM layer-order relation is part of what is pinned."""
    import os
    import numpy as np
    from oracle import rcwa_oracle as O
    g = np.load(os.path.join(golden_dir, "rcwa_stack_golden.npz"))
    names = sorted({k.split("/")[0] for k in g.files})
    assert len(names) == 10
    for n in names:
        r = O.evaluate_stack(g[n + "/structs"], float(g[n + "/wavelength"]), float(g[n + "/angle"]), int(g[n + "/order"]),
                             g[n + "/thickness"], int(g[n + "/pol"]), str(g[n + "/algo"]))
        assert abs(r.eff - float(g[n + "/ref_eff"])) <= 1e-12, n
        assert np.abs(r.tran_cut - g[n + "/ref_tran_cut"]).max() <= 1e-12 and np.abs(r.refl_cut - g[n + "/ref_refl_cut"]).max() <= 1e-12
    # the quirk itself: SMM(list) == TMM(reversed list) for a two-layer stack
    s = g["ML2_block16_N20_TMM/structs"]; th = g["ML2_block16_N20_TMM/thickness"]
    a = O.evaluate_stack(s, 1100.0, 60.0, 20, th, 1, 'SMM').eff
    b = O.evaluate_stack(s[::-1], 1100.0, 60.0, 20, th[::-1], 1, 'TMM').eff
    assert abs(a - b) <= 1e-11 and abs(a - float(g["ML2_block16_N20_SMM/ref_eff"])) <= 1e-12


def test_oracle_smm_against_sixteen_reference_smm_runs(golden_dir):
    """The oracle's S-matrix form (solve_stack, algo='SMM') against 16 outputs of the unmodified reference run with
    algo='SMM' at order 60 (oracle/make_smm_golden.py): blocky and random patterns, mixed wavelength / angle / thickness."""
    import os
    import numpy as np
    from oracle import rcwa_oracle as O
    g = np.load(os.path.join(golden_dir, "rcwa_smm_golden.npz"))
    order = int(g["order"])
    for i in range(len(g["ref_eff"])):
        r = O.evaluate_stack(g["patterns"][i][None, :], float(g["wavelength"][i]), float(g["angle"][i]), order,
                             (float(g["thickness"][i]),), 1, 'SMM')
        assert abs(r.eff - g["ref_eff"][i]) <= 1e-12
        assert np.abs(r.tran_cut - g["ref_tran_cut"][i]).max() <= 1e-12 and np.abs(r.refl_cut - g["ref_refl_cut"][i]).max() <= 1e-12
