"""GPU (B200): the CUDA path, called through the C-ABI library, against

  * the committed fixtures produced by the UNMODIFIED reference solver (tests/golden/*.npz, oracle/make_golden.py),
  * the CPU oracle (oracle/rcwa_oracle.py) on freshly seeded inputs at sizes it finishes in seconds,
  * size-independent properties at BASELINE.json's full batch sizes (translation / mirror symmetry, chunking
    idempotence, determinism).

Tolerance: |d efficiency| <= 1e-6 absolute (north_star, complex128 mode).  The observed error is ~1e-12; a second,
tighter assertion (1e-9) guards against silent precision regressions.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-6        # north_star: diffraction efficiencies within 1e-6 absolute in complex128
TIGHT = 1e-9      # regression guard (float64 path; measured ~1e-12)


@pytest.fixture(scope="module")
def torch_mod():
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    return torch


@pytest.fixture(scope="module")
def pkg():
    import physics_informed_reinforcement_learning_b200 as p
    return p


def _named(golden_dir):
    g = np.load(os.path.join(golden_dir, "rcwa_golden.npz"))
    names = sorted({k.split("/")[0] for k in g.files})
    return g, names


def test_named_golden_cases_through_c_abi_host_entry(golden_dir, pkg, torch_mod):
    """Every named reference case (G1..G8, Reticolo known answer), through pirl_rcwa_eval_host."""
    g, names = _named(golden_dir)
    backends = {}
    worst = 0.0
    for n in names:
        pat = g[n + "/pattern"]
        order = int(g[n + "/order"])
        key = (pat.size, order)
        if key not in backends:
            backends[key] = pkg.RCWABatch(pat.size, order, max_batch=8)
        be = backends[key]
        eff, r3, t3, st = be.evaluate_host(pat[None, :], float(g[n + "/wavelength"]), float(g[n + "/angle"]),
                                           float(g[n + "/thickness"]), return_orders=True)
        assert st[0] == 0, (n, st)
        e = abs(eff[0] - float(g[n + "/ref_eff"]))
        # reference cuts are [::-1] of de[N-1:N+2]
        et = np.abs(t3[0][::-1] - g[n + "/ref_tran_cut"]).max()
        er = np.abs(r3[0][::-1] - g[n + "/ref_refl_cut"]).max()
        worst = max(worst, e, et, er)
        assert e <= TOL and et <= TOL and er <= TOL, (n, e, et, er)
    assert worst <= TIGHT, worst
    for be in backends.values():
        be.close()


def test_reticolo_known_answer(golden_dir, pkg):
    g, _ = _named(golden_dir)
    n = "KA_reticolo_800nm_70deg"
    be = pkg.RCWABatch(256, 40, max_batch=4)
    eff = be.evaluate_host(np.stack([g[n + "/pattern"], g[n + "/pattern"][::-1]]), 800.0, 70.0, 325.0)
    assert abs(eff[0] - float(g[n + "/reticolo_efficiency"])) < 5e-8      # float32 rounding of the recorded value
    assert abs(eff[1] - 0.027283368919125597) < TIGHT                     # orientation pin
    be.close()


def test_batch_fixture_device_entry(golden_dir, pkg, torch_mod):
    """256 mixed random / blocky patterns at the headline geometry, device-resident entry (pirl_rcwa_eval)."""
    torch = torch_mod
    b = np.load(os.path.join(golden_dir, "rcwa_batch.npz"))
    be = pkg.RCWABatch(256, 40, max_batch=100)        # 256 = 100 + 100 + 56: exercises ragged chunking
    pats = torch.from_numpy(b["patterns"]).cuda()
    eff, r3, t3, st = be.evaluate(pats, 1100.0, 60.0, 325.0, return_orders=True)
    torch.cuda.synchronize()
    assert int(st.abs().sum()) == 0
    eff, r3, t3 = eff.cpu().numpy(), r3.cpu().numpy(), t3.cpu().numpy()
    assert np.abs(eff - b["ref_eff"]).max() <= TIGHT
    assert np.abs(t3[:, ::-1] - b["ref_tran_cut"]).max() <= TIGHT
    assert np.abs(r3[:, ::-1] - b["ref_refl_cut"]).max() <= TIGHT
    assert b["ref_eff"].max() > 0.5
    be.close()


def test_sweep_fixture_per_structure_parameters(golden_dir, pkg, torch_mod):
    """(wavelength, angle) grid at N=60 evaluated as ONE batch with per-structure parameter rows (config C4)."""
    s = np.load(os.path.join(golden_dir, "rcwa_sweep.npz"))
    grid = s["grid"]
    pats = s["patterns"][grid[:, 2].astype(int)]
    be = pkg.RCWABatch(256, int(s["order"]), max_batch=64)
    eff = be.evaluate_host(pats, grid[:, 0], grid[:, 1], float(s["thickness"]))
    assert np.abs(eff - grid[:, 3]).max() <= TIGHT
    be.close()


@pytest.mark.parametrize("n_cells,order,count", [(256, 40, 48), (64, 8, 32), (256, 60, 12), (512, 100, 3), (100, 17, 8), (128, 45, 6), (128, 47, 4), (128, 41, 4)])
def test_against_oracle_on_seeded_inputs(pkg, n_cells, order, count):
    from oracle import rcwa_oracle as O
    rng = np.random.default_rng(1000 + n_cells + order)
    pats = rng.choice([-1, 1], (count, n_cells)).astype(np.int8)
    for i in range(1, count, 3):                          # blocky rows: efficiencies far from the noise floor
        blk = (4, 10, 16)[i % 3] if n_cells % 16 else (8, 16, 32)[i % 3]
        base = rng.choice([-1, 1], -(-n_cells // blk))
        pats[i] = np.repeat(base, blk)[:n_cells]
    wl, ang, thick = 1100.0, 60.0, 325.0
    be = pkg.RCWABatch(n_cells, order, max_batch=16)
    eff, r3, t3, st = be.evaluate_host(pats, wl, ang, thick, return_orders=True)
    assert not st.any()
    for i in range(count):
        r = O.evaluate(pats[i].astype(np.float64), wl, ang, order, thick)
        assert abs(eff[i] - r.eff) <= TIGHT, (i, eff[i], r.eff)
        assert np.abs(t3[i] - r.de_ti[order - 1:order + 2]).max() <= TIGHT
        assert np.abs(r3[i] - r.de_ri[order - 1:order + 2]).max() <= TIGHT
    be.close()


def test_every_order_against_oracle(pkg):
    """One handle per order (every order up to 70, every third up to the maximum 126): crosses every kernel-variant boundary of
    the launch plans (k_apply row / sweep variants at ff = 81 / 96, k_tail column variants at 96 / 128, the register-tile sizes
    81 and 121) and, with 96 cells, the orders whose harmonics alias (2 * order >= n_cells), two structures each."""
    from oracle import rcwa_oracle as O
    rng = np.random.default_rng(4242)
    n_cells, wl, ang, thick = 96, 1050.0, 58.0, 310.0
    worst = 0.0
    for order in list(range(1, 71)) + list(range(72, 127, 3)):
        pats = np.stack([rng.choice([-1, 1], n_cells), np.repeat(rng.choice([-1, 1], n_cells // 8), 8)]).astype(np.int8)
        be = pkg.RCWABatch(n_cells, order, max_batch=4)
        eff, st = be.evaluate_host(pats, wl, ang, thick), None
        for i in range(2):
            r = O.evaluate(pats[i].astype(np.float64), wl, ang, order, thick)
            worst = max(worst, abs(eff[i] - r.eff))
            assert abs(eff[i] - r.eff) <= TIGHT, (order, i, eff[i], r.eff)
        be.close()


@pytest.mark.parametrize("n_cells", [2, 3, 7, 255, 257, 1000, 20000, 65536])
def test_cell_counts_from_two_to_the_maximum(pkg, n_cells):
    """Odd, tiny and very long patterns at the headline order; beyond ~10^4 cells the register-tile front kernel no longer
    holds the pattern tables in shared memory and the handle plans the generic kernels instead."""
    from oracle import rcwa_oracle as O
    rng = np.random.default_rng(n_cells)
    pats = np.stack([rng.choice([-1, 1], n_cells), np.repeat(rng.choice([-1, 1], -(-n_cells // 5)), 5)[:n_cells]]).astype(np.int8)
    be = pkg.RCWABatch(n_cells, 40, max_batch=4)
    eff, r3, t3, st = be.evaluate_host(pats, 1100.0, 60.0, 325.0, return_orders=True)
    assert not st.any()
    for i in range(2):
        r = O.evaluate(pats[i].astype(np.float64), 1100.0, 60.0, 40, 325.0)
        assert abs(eff[i] - r.eff) <= TIGHT, (n_cells, i, eff[i], r.eff)
    be.close()


def test_extreme_thickness_wavelength_angle(pkg):
    from oracle import rcwa_oracle as O
    pats = np.repeat(np.random.default_rng(5).choice([-1, 1], (1, 32)), 8, axis=1).astype(np.int8)
    be = pkg.RCWABatch(256, 40, max_batch=4)
    for thick in (1.0, 100.0, 5000.0, 50000.0):
        for wl, ang in ((1100.0, 60.0), (400.0, 20.0), (2000.0, 85.0), (1100.0, 5.0)):
            eff, st = be.evaluate_host(pats, wl, ang, thick), None
            ref = O.evaluate(pats[0].astype(np.float64), wl, ang, 40, thick).eff
            assert abs(eff[0] - ref) <= TIGHT, (thick, wl, ang, eff[0], ref)
    be.close()


def test_other_wavelengths_angles_thickness_vs_oracle(pkg):
    from oracle import rcwa_oracle as O
    rng = np.random.default_rng(77)
    pats = np.repeat(rng.choice([-1, 1], (6, 32)), 8, axis=1).astype(np.int8)
    wls = np.array([900.0, 950.0, 1000.0, 1100.0, 1200.0, 800.0])
    angs = np.array([50.0, 55.0, 65.0, 70.0, 75.0, 60.0])
    ths = np.array([325.0, 300.0, 250.0, 400.0, 325.0, 500.0])
    be = pkg.RCWABatch(256, 40, max_batch=8)
    eff = be.evaluate_host(pats, wls, angs, ths)
    ref = np.array([O.efficiency(p.astype(float), w, a, 40, t) for p, w, a, t in zip(pats, wls, angs, ths)])
    assert np.abs(eff - ref).max() <= TIGHT
    be.close()


def test_trivial_and_edge_inputs(pkg, torch_mod):
    torch = torch_mod
    be = pkg.RCWABatch(256, 40, max_batch=8)
    pats = np.ones((3, 256), dtype=np.int8)
    pats[1] = -1
    pats[2, ::2] = -1                                        # finest possible grating
    eff, r3, t3, st = be.evaluate_host(pats, return_orders=True)
    assert not st.any()
    assert abs(eff[0]) < 1e-15 and abs(eff[1]) < 1e-15        # uniform slabs diffract nothing (G3)
    assert abs(r3[0, 1] + t3[0, 1] - 1.0) < 1e-10             # ... and conserve energy in the zeroth order
    # empty batch: allowed, returns empty
    e0 = be.evaluate(torch.empty(0, 256, dtype=torch.int8, device="cuda"))
    assert e0.numel() == 0
    assert be.evaluate_host(np.ones((0, 256))).size == 0
    # wrong shapes / dtypes are rejected on the host
    with pytest.raises(ValueError):
        be.evaluate_host(np.ones((2, 255)))
    with pytest.raises(TypeError):
        be.evaluate(torch.ones(2, 256, device="cuda"))
    with pytest.raises(ValueError):
        be.evaluate_host(np.ones((3, 256)), wavelength=[1000.0, 1100.0])
    be.close()


def test_full_batch_properties_4096(pkg, torch_mod):
    """BASELINE config C2 size (4096 x 256, N=40): properties that need no oracle.

    * a circular shift of the pattern only multiplies every Fourier coefficient by a phase -> efficiencies unchanged;
    * mirroring x -> -x swaps the +1 and -1 orders;
    * chunked evaluation == one-shot evaluation, bit for bit; repeated evaluation is deterministic;
    * lossless structure: de_r, de_t in [0, 1], the three central orders never exceed total power."""
    torch = torch_mod
    rng = np.random.default_rng(1234)
    pats = (rng.integers(0, 2, (4096, 256)) * 2 - 1).astype(np.int8)
    pats[::4] = np.repeat(rng.choice([-1, 1], (1024, 16)), 16, axis=1)
    be = pkg.RCWABatch(256, 40, max_batch=4096)
    dp = torch.from_numpy(pats).cuda()
    eff, r3, t3, st = be.evaluate(dp, return_orders=True)
    eff2 = be.evaluate(dp)
    shifts = torch.from_numpy(rng.integers(1, 256, 4096)).cuda()
    idx = (torch.arange(256, device="cuda")[None, :] + shifts[:, None]) % 256
    eff_shift = be.evaluate(torch.gather(dp, 1, idx))
    effm, r3m, t3m, _ = be.evaluate(torch.flip(dp, dims=[1]).contiguous(), return_orders=True)
    be_small = pkg.RCWABatch(256, 40, max_batch=1000)
    eff_chunk = be_small.evaluate(dp)
    torch.cuda.synchronize()
    assert int(st.abs().sum()) == 0
    assert torch.equal(eff, eff2)
    assert torch.equal(eff, eff_chunk)
    assert (eff - eff_shift).abs().max().item() <= 1e-9
    assert (t3[:, 0] - t3m[:, 2]).abs().max().item() <= 1e-9 and (t3[:, 2] - t3m[:, 0]).abs().max().item() <= 1e-9
    assert (r3[:, 0] - r3m[:, 2]).abs().max().item() <= 1e-9
    tot = r3.sum(1) + t3.sum(1)
    assert (tot <= 1.0 + 1e-9).all() and (r3 >= -1e-12).all() and (t3 >= -1e-12).all()
    assert eff.max().item() > 0.3
    be.close(); be_small.close()


def test_full_batch_sample_vs_oracle_4096(pkg, torch_mod):
    """The C2 workload itself (default_rng(1234) patterns): a strided sample of 64 of the 4096 against the oracle."""
    from oracle import rcwa_oracle as O
    torch = torch_mod
    pats = (np.random.default_rng(1234).integers(0, 2, (4096, 256)) * 2 - 1).astype(np.int8)
    be = pkg.RCWABatch(256, 40, max_batch=4096)
    eff = be.evaluate(torch.from_numpy(pats).cuda()).cpu().numpy()
    idx = np.arange(0, 4096, 64)
    ref = np.array([O.efficiency(pats[i].astype(float)) for i in idx])
    assert np.abs(eff[idx] - ref).max() <= TIGHT
    be.close()


def test_jlabcode_and_meentindex_drop_in(golden_dir, pkg):
    """The reference-facing API: JLABCode(...).reproduce_acs_cell and MeentIndex.reset/step/get_efficiency."""
    g, _ = _named(golden_dir)
    n = "G7_block16_N40"
    pat = g[n + "/pattern"]
    wls = np.array([1100])
    period = abs(wls / np.sin(60 / 180 * np.pi))
    calc = pkg.JLABCode(grating_type=0, n_I=1.45, n_II=1., theta=0, phi=0., fourier_order=40, period=period, wls=wls,
                        pol=1, patterns=None, ucell=pat[np.newaxis, np.newaxis, :], thickness=np.array([325]))
    eff, refl_cut, tran_cut = calc.reproduce_acs_cell('p_si__real', 1)
    assert abs(eff - float(g[n + "/ref_eff"])) <= TIGHT
    assert np.abs(tran_cut - g[n + "/ref_tran_cut"]).max() <= TIGHT
    assert np.abs(refl_cut - g[n + "/ref_refl_cut"]).max() <= TIGHT
    with pytest.raises(ValueError):
        pkg.JLABCode(grating_type=1, ucell=pat[None, None, :])
    with pytest.raises(ValueError):
        pkg.JLABCode(pol=2, ucell=pat[None, None, :])

    env = pkg.make('MeentIndex-v0', wavelength=1100, desired_angle=60, thickness=325)
    assert env.observation_space.shape == (256,) and env.action_space.n == 256
    np.random.seed(42)
    obs = env.reset()
    assert list(np.nonzero(obs == -1)[0]) == [31, 85, 176, 253]                    # SURVEY G5 (bit-exact reset)
    assert abs(env.eff - float(g["G5_bunch_seed42/ref_eff"])) <= TIGHT
    e0 = env.eff
    obs2, reward, done, info = env.step(100)
    assert obs2[100] == -1 and done is False and info == {}
    assert reward == env.eff - e0
    obs3, r3, _, _ = env.step(256)                                                 # out of range: no-op
    assert np.array_equal(obs3, obs2) and r3 == 0.0
    # G8: flip delta on the block pattern
    env.struct = pat.copy()
    env.eff = env.get_efficiency(env.struct)
    _, rew, _, _ = env.step(100)
    ref_delta = float(g["G8_block16_flip100/ref_eff"]) - float(g[n + "/ref_eff"])
    assert abs(rew - ref_delta) <= 2 * TIGHT


def test_vector_env_matches_scalar_semantics_bit_exact(pkg, torch_mod):
    """VectorDeflectorEnv vs the restated MeentIndex semantics (oracle/env_oracle.py) driven by the SAME efficiency
    values: patterns, observations, dones bit-exact; rewards equal to the float64 difference of the two efficiencies."""
    from oracle import env_oracle
    torch = torch_mod
    B, T = 6, 9
    be = pkg.RCWABatch(256, 40, max_batch=64)
    env = pkg.VectorDeflectorEnv(B, backend=be, max_episode_steps=4)
    np.random.seed(7)
    obs = env.reset()
    np.random.seed(7)
    scal = [env_oracle.ScalarEnv(lambda s: float(be.evaluate_host(s[None, :])[0]), max_episode_steps=4) for _ in range(B)]
    sobs = np.stack([e.reset() for e in scal])
    assert np.array_equal(obs.cpu().numpy(), sobs)
    assert np.array_equal(env.eff.cpu().numpy(), np.array([e.eff for e in scal]))
    rng = np.random.default_rng(5)
    for t in range(T):
        acts = rng.integers(-2, 260, B)
        obs, rew, done, info = env.step(acts, auto_reset=False)
        outs = [e.step(int(a)) for e, a in zip(scal, acts)]
        assert np.array_equal(obs.cpu().numpy(), np.stack([o[0] for o in outs]))
        assert np.array_equal(rew.cpu().numpy(), np.array([o[1] for o in outs]))      # bit-exact f64 subtraction
        assert np.array_equal(done.cpu().numpy(), np.array([o[2] for o in outs]))
        assert np.array_equal(env.best_eff.cpu().numpy(), np.array([e.best[0] for e in scal]))
        assert np.array_equal(env.best_patterns.cpu().numpy(), np.stack([e.best[1] for e in scal]).astype(np.int8))
        if done.any():
            state = np.random.get_state()            # both sides must consume the SAME global-RNG draws
            env.reset(np.nonzero(done.cpu().numpy())[0])
            np.random.set_state(state)
            for e, d in zip(scal, done.cpu().numpy()):
                if d:
                    e.reset()
            assert np.array_equal(env.obs().cpu().numpy(), np.stack([e.struct for e in scal]))
    be.close()


def test_stage_timers_and_launch_counter(pkg, torch_mod):
    torch = torch_mod
    be = pkg.RCWABatch(256, 40, max_batch=32)
    pats = torch.from_numpy(np.random.default_rng(0).choice([-1, 1], (64, 256)).astype(np.int8)).cuda()
    n0 = be.launch_count()
    be.set_stage_timing(True)
    be.evaluate(pats)
    times = be.stage_times()
    be.set_stage_timing(False)
    launched = be.launch_count() - n0
    assert launched == sum(c for _, c in times.values()) > 0
    assert all(ms > 0 for ms, c in times.values() if c)
    be.close()


def test_register_tile_paths_against_shared_memory_kernels_and_pipeline_settings(pkg, torch_mod):
    """Orders 40 and 60 run register-tile kernels (order 60 on a padded 126-tile grid) and, for large chunks, the sub-chunk
    pipeline.  The generic shared-memory kernels (PIRL_NO_REGTILE=1) are an independent implementation of the same path: both
    must agree far below the parity tolerance, and the launch order (pipeline on / off / other split) must not change a bit."""
    torch = torch_mod
    rng = np.random.default_rng(77)
    for order, count in ((40, 1300), (60, 700)):
        pats = (rng.integers(0, 2, (count, 256)) * 2 - 1).astype(np.int8)
        pats[:64] = np.repeat(rng.integers(0, 2, (64, 16)) * 2 - 1, 16, axis=1)
        d = torch.from_numpy(pats).cuda()
        be = pkg.RCWABatch(256, order, max_batch=count, device=0)
        e_auto, r3, t3, st = be.evaluate(d, return_orders=True)
        be.set_subchunk(0)
        e_off = be.evaluate(d)
        be.set_subchunk(300)
        e_300 = be.evaluate(d)
        torch.cuda.synchronize()
        assert int(st.abs().sum()) == 0
        assert torch.equal(e_auto, e_off) and torch.equal(e_auto, e_300)
        os.environ["PIRL_NO_REGTILE"] = "1"
        try:
            ref = pkg.RCWABatch(256, order, max_batch=256, device=0)
        finally:
            del os.environ["PIRL_NO_REGTILE"]
        e_ref, r3_ref, t3_ref, _ = ref.evaluate(d[:256], return_orders=True)
        torch.cuda.synchronize()
        assert float((e_ref - e_auto[:256]).abs().max()) <= TIGHT
        assert float((r3_ref - r3[:256]).abs().max()) <= TIGHT and float((t3_ref - t3[:256]).abs().max()) <= TIGHT
        # lossless structure: the three reported orders never exceed unity and the efficiencies are non-negative
        assert float(e_auto.min()) >= 0.0 and float((r3.sum(1) + t3.sum(1)).max()) <= 1.0 + 1e-9
        be.close(); ref.close()


def _oracle_pool_eval(pats, wl, ang, order, thick):
    """All host cores: the order-100 oracle costs 0.1-0.2 s per structure."""
    from oracle import cpu_pool
    pool = cpu_pool.OraclePool()
    try:
        eff, _ = pool.run(pats, wl, ang, order, thick)
    finally:
        pool.close()
    return eff


def test_config_c5_order100_256_patterns_vs_oracle(pkg, torch_mod):
    """BASELINE config 5 (order 100, ff = 201, 512 cells) on the thread-block-cluster kernels: 256 patterns -- random, and
    8- / 16- / 32-pixel blocks so that efficiencies span 0 ... 0.9 -- against the oracle, all three reported orders of a
    subset, and the same batch through the generic shared-memory kernels (PIRL_NO_CLUSTER=1), an independent implementation."""
    torch = torch_mod
    from oracle import rcwa_oracle as O
    rng = np.random.default_rng(5)
    n, order, B = 512, 100, 256
    pats = (rng.integers(0, 2, (B, n)) * 2 - 1).astype(np.int8)
    for i, blk in ((1, 8), (2, 16), (3, 32)):
        pats[i::4] = np.repeat(rng.integers(0, 2, (B // 4, n // blk)) * 2 - 1, blk, axis=1)
    be = pkg.RCWABatch(n, order, max_batch=B, device=0)
    be.set_stage_timing(True)
    eff, r3, t3, st = be.evaluate(torch.from_numpy(pats).cuda(), return_orders=True)
    torch.cuda.synchronize()
    times = be.stage_times()
    be.set_stage_timing(False)
    assert int(st.abs().sum()) == 0
    assert times["ck_front"][1] == 1 and times["ck_tail"][1] == 1 and times["k_front"][1] == 0     # the cluster kernels ran
    eff, r3, t3 = eff.cpu().numpy(), r3.cpu().numpy(), t3.cpu().numpy()
    ref = _oracle_pool_eval(pats, 1100.0, 60.0, order, 325.0)
    err = np.abs(eff - ref)
    assert err.max() <= TOL, (int(err.argmax()), err.max())
    assert err.max() <= 1e-8                           # observed 1e-9 class; the contract is 1e-6
    assert ref.max() > 0.3                             # the blocky rows reach large efficiencies
    for i in (0, 1, 2, 3, 77, 130, 255):
        r = O.evaluate(pats[i].astype(np.float64), 1100.0, 60.0, order, 325.0)
        assert np.abs(t3[i] - r.de_ti[order - 1:order + 2]).max() <= TOL and np.abs(r3[i] - r.de_ri[order - 1:order + 2]).max() <= TOL
    os.environ["PIRL_NO_CLUSTER"] = "1"
    try:
        gen = pkg.RCWABatch(n, order, max_batch=32, device=0)
    finally:
        del os.environ["PIRL_NO_CLUSTER"]
    e_gen = gen.evaluate_host(pats[:32])
    assert np.abs(e_gen - eff[:32]).max() <= 1e-8
    be.close(); gen.close()


@pytest.mark.parametrize("n_cells,order", [(256, 50), (256, 61), (300, 77), (512, 88), (512, 103)])
def test_cluster_path_orders_vs_oracle(pkg, n_cells, order):
    """Orders 41-59 and 61-103 run the cluster kernels on the identity-padded 208 x 208 tile grid: other paddings than order 100."""
    rng = np.random.default_rng(order)
    pats = (rng.integers(0, 2, (24, n_cells)) * 2 - 1).astype(np.int8)
    pats[::3] = np.repeat(rng.integers(0, 2, (8, -(-n_cells // 16))) * 2 - 1, 16, axis=1)[:, :n_cells]
    be = pkg.RCWABatch(n_cells, order, max_batch=16, device=0)        # 24 > 16: two chunks
    wl, ang, thick = 1000.0, 65.0, 300.0
    eff, r3, t3, st = be.evaluate_host(pats, wl, ang, thick, return_orders=True)
    assert not st.any()
    ref = _oracle_pool_eval(pats, wl, ang, order, thick)
    assert np.abs(eff - ref).max() <= 1e-8, np.abs(eff - ref).max()
    assert (r3.sum(1) + t3.sum(1)).max() <= 1.0 + 1e-9
    be.close()


def test_cluster_small_variant_matches_register_tile_kernels(pkg, torch_mod):
    """PIRL_FORCE_CLUSTER=1 routes order 40 through the 84 x 84 instance of the cluster kernels (same code, four CTAs, small
    tiles): against the single-SM register-tile kernels on 600 structures, work queue and ragged last wave included."""
    torch = torch_mod
    rng = np.random.default_rng(8)
    pats = torch.from_numpy((rng.integers(0, 2, (600, 256)) * 2 - 1).astype(np.int8)).cuda()
    ref = pkg.RCWABatch(256, 40, max_batch=600, device=0)
    e_ref, r_ref, t_ref, _ = ref.evaluate(pats, return_orders=True)
    os.environ["PIRL_FORCE_CLUSTER"] = "1"
    os.environ["PIRL_NO_REGTILE"] = "1"
    try:
        ck = pkg.RCWABatch(256, 40, max_batch=600, device=0)
    finally:
        del os.environ["PIRL_FORCE_CLUSTER"], os.environ["PIRL_NO_REGTILE"]
    ck.set_stage_timing(True)
    e, r, t, st = ck.evaluate(pats, return_orders=True)
    torch.cuda.synchronize()
    times = ck.stage_times()
    assert int(st.abs().sum()) == 0 and times["ck_front"][1] == 1 and times["ck_tail"][1] == 1
    assert float((e - e_ref).abs().max()) <= 1e-9 and float((t - t_ref).abs().max()) <= 1e-9 and float((r - r_ref).abs().max()) <= 1e-9
    ref.close(); ck.close()


def test_smm_scattering_matrix_operator(golden_dir, pkg, torch_mod):
    """algo = 'SMM' (scattering_method.py:12-73, smm_util.py): k_smm builds S_ref, S_layer, S_trn and combines them with
    Redheffer star products.  Against the reference's own SMM result (golden G1_rand256_N40_SMM), against the oracle on
    random / blocky patterns at three orders (ETM and SMM agree to ~1e-13 in the reference), against
    the ETM kernels of this library, and through the JLABCode drop-in."""
    torch = torch_mod
    from oracle import rcwa_oracle as O
    rng = np.random.default_rng(21)
    for n_cells, order, count in ((256, 40, 24), (96, 12, 16), (256, 64, 6)):
        pats = (rng.integers(0, 2, (count, n_cells)) * 2 - 1).astype(np.int8)
        pats[::2] = np.repeat(rng.integers(0, 2, (count // 2, n_cells // 16)) * 2 - 1, 16, axis=1)
        smm = pkg.RCWABatch(n_cells, order, max_batch=16, device=0, algo='SMM')
        etm = pkg.RCWABatch(n_cells, order, max_batch=16, device=0)
        smm.set_stage_timing(True)
        e_s, r_s, t_s, st = smm.evaluate_host(pats, 1100.0, 60.0, 325.0, return_orders=True)
        times = smm.stage_times()
        e_e, r_e, t_e, _ = etm.evaluate_host(pats, 1100.0, 60.0, 325.0, return_orders=True)
        assert not st.any() and times["k_smm"][1] >= 1 and times["k_tail"][1] == 0 and times["rk_tail"][1] == 0
        assert np.abs(e_s - e_e).max() <= 1e-9 and np.abs(r_s - r_e).max() <= 1e-9 and np.abs(t_s - t_e).max() <= 1e-9
        for i in range(0, count, 3):
            r = O.evaluate(pats[i].astype(np.float64), 1100.0, 60.0, order, 325.0)
            assert abs(e_s[i] - r.eff) <= 1e-9
            assert np.abs(t_s[i] - r.de_ti[order - 1:order + 2]).max() <= 1e-9 and np.abs(r_s[i] - r.de_ri[order - 1:order + 2]).max() <= 1e-9
        with pytest.raises(ValueError):
            smm.evaluate_grad(torch.from_numpy(pats).cuda())            # the adjoint sweep exists for the ETM form only
        smm.close(); etm.close()
    # the reference's own SMM run of G1 (fixture written by oracle/make_golden.py from the unmodified reference, algo='SMM')
    g, names = _named(golden_dir)
    assert "G1_rand256_N40_SMM" in names
    pat = g["G1_rand256_N40_SMM/pattern"]
    assert abs(float(g["G1_rand256_N40_SMM/ref_eff"]) - 0.0004665720258772289) <= 1e-15
    code = pkg.JLABCode(n_I=1.45, n_II=1., theta=0, phi=0, fourier_order=40, period=[abs(1100 / np.sin(60 / 180 * np.pi))],
                        wls=np.array([1100.]), pol=1, ucell=pat[None, None, :], thickness=[325], algo='SMM')
    eff, refl, tran = code.reproduce_acs_cell('p_si__real', 1)
    assert abs(eff - float(g["G1_rand256_N40_SMM/ref_eff"])) <= 1e-12
    assert np.abs(tran - g["G1_rand256_N40_SMM/ref_tran_cut"]).max() <= 1e-12 and np.abs(refl - g["G1_rand256_N40_SMM/ref_refl_cut"]).max() <= 1e-12


def test_layer_stacks_and_te_against_the_unmodified_reference(golden_dir, pkg):
    """SURVEY section 8f rank 4, partly: multilayer stacks (_base.py:141) and TE polarisation (_base.py:142-148) through
    pirl_rcwa_eval_stack / the TE front + k_smm, against fixtures written by oracle/make_stack_golden.py from the unmodified
    reference (JLABCode with ucell of shape (L, 1, n)), for 'TMM' and 'SMM' -- including the reference's layer-order quirk
    (its SMM result of a list equals its TMM result of the reversed list) -- through RCWABatch and the JLABCode drop-in."""
    g = np.load(os.path.join(golden_dir, "rcwa_stack_golden.npz"))
    names = sorted({k.split("/")[0] for k in g.files})
    assert len(names) == 10
    worst = 0.0
    for n in names:
        structs = g[n + "/structs"]
        wl, ang, order = float(g[n + "/wavelength"]), float(g[n + "/angle"]), int(g[n + "/order"])
        th, pol, algo = g[n + "/thickness"], int(g[n + "/pol"]), str(g[n + "/algo"])
        be = pkg.RCWABatch(structs.shape[1], order, pol=pol, max_batch=16, device=0, algo=algo)
        eff, r3, t3, st = be.evaluate_stack_host(structs[None], wl, ang, th, return_orders=True)
        assert st[0] == 0, (n, st)
        e = max(abs(eff[0] - float(g[n + "/ref_eff"])), np.abs(t3[0][::-1] - g[n + "/ref_tran_cut"]).max(),
                np.abs(r3[0][::-1] - g[n + "/ref_refl_cut"]).max())
        worst = max(worst, e)
        assert e <= TIGHT, (n, e)
        if structs.shape[0] == 1:            # single layer: the plain entry point must agree with the stack entry point
            e1 = be.evaluate_host(structs[0][None], wl, ang, float(th[0]))
            assert abs(e1[0] - eff[0]) <= 1e-12
        code = pkg.JLABCode(n_I=1.45, n_II=1., theta=0, phi=0, fourier_order=order, period=[abs(wl / np.sin(ang / 180 * np.pi))],
                            wls=np.array([wl]), pol=pol, ucell=structs[:, None, :], thickness=th, algo=algo)
        ej, refl, tran = code.reproduce_acs_cell('p_si__real', 1)
        assert abs(ej - float(g[n + "/ref_eff"])) <= TIGHT and np.abs(tran - g[n + "/ref_tran_cut"]).max() <= TIGHT
        be.close()
    # batch of stacks with per-structure thicknesses vs the oracle restatement (pinned to the reference above)
    from oracle import rcwa_oracle as O
    rng = np.random.default_rng(9)
    pats = np.repeat(rng.choice([-1, 1], (5, 3, 16)), 16, axis=2)
    ths = rng.uniform(60.0, 300.0, (5, 3))
    be = pkg.RCWABatch(256, 24, max_batch=8, device=0)              # 15 layer problems > max_batch: chunked by structure
    eff, r3, t3, st = be.evaluate_stack_host(pats, 1050.0, 55.0, ths, return_orders=True)
    assert not st.any()
    for i in range(5):
        r = O.evaluate_stack(pats[i].astype(float), 1050.0, 55.0, 24, ths[i], 1, 'TMM')
        assert abs(eff[i] - r.eff) <= TIGHT and np.abs(t3[i] - r.de_ti[23:26]).max() <= TIGHT
    be.close()


def test_device_and_host_entry_points_share_one_handle_safely(pkg, torch_mod):
    """evaluate() is asynchronous on the caller's stream, evaluate_host() runs on the handle's own stream, and both use the
    handle's workspaces: a host call issued while a device call is still running must wait for it (and the other way round)."""
    torch = torch_mod
    rng = np.random.default_rng(11)
    be = pkg.RCWABatch(256, 40, max_batch=1024)
    big = torch.from_numpy(rng.choice(np.array([-1, 1], dtype=np.int8), size=(1024, 256))).cuda()
    small = rng.choice(np.array([-1, 1], dtype=np.int8), size=(3, 256))
    want_big = be.evaluate(big).clone()
    torch.cuda.synchronize()
    want_small = be.evaluate_host(small).copy()
    side = torch.cuda.Stream()
    for _ in range(5):
        got_big = be.evaluate(big)                       # ~2 ms of kernels in flight on the current stream ...
        got_small = be.evaluate_host(small)              # ... while this call starts on the handle's own stream
        with torch.cuda.stream(side):
            got_side = be.evaluate(big[:64])             # and a third stream right behind it
        torch.cuda.synchronize()
        assert torch.equal(got_big, want_big)
        assert np.array_equal(got_small, want_small)
        assert torch.equal(got_side, want_big[:64])
    be.close()


def test_env_kernels_reproduce_reference_step_fixtures(golden_dir, torch_mod):
    """pirl_env_flip / pirl_env_reward through the C ABI against trajectories produced by the reference's own flip / step
    method objects (oracle/make_env_step_golden.py): every (state, action) -> state bit-exact, out-of-range actions
    included; reward = eff_new - eff_prev bit-exact in float64 from the recorded efficiencies."""
    import ctypes as C
    from physics_informed_reinforcement_learning_b200 import _lib as lib_mod
    torch = torch_mod
    lib = lib_mod.load()
    g = np.load(os.path.join(golden_dir, "env_step_golden.npz"))
    K, T = g["actions"].shape
    pats = torch.from_numpy(g["init"].astype(np.int8)).cuda()                 # the K trajectories advance as one batch
    prev = torch.from_numpy(g["effs"][:, 0].copy()).cuda()
    rew = torch.empty(K, dtype=torch.float64, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for t in range(T):
        acts = torch.from_numpy(g["actions"][:, t].astype(np.int64)).cuda()
        lib_mod.check(lib.pirl_env_flip(pats.data_ptr(), acts.data_ptr(), K, 256, st))
        new = torch.from_numpy(g["effs"][:, t + 1].copy()).cuda()
        lib_mod.check(lib.pirl_env_reward(new.data_ptr(), prev.data_ptr(), rew.data_ptr(), None, None, pats.data_ptr(), K, 256, st))
        torch.cuda.synchronize()
        assert np.array_equal(pats.cpu().numpy(), g["states"][:, t].astype(np.int8)), t
        assert np.array_equal(rew.cpu().numpy(), g["rewards"][:, t]), t
        assert np.array_equal(prev.cpu().numpy(), g["effs"][:, t + 1]), t      # eff_prev <- eff_new


def test_smm_operator_against_sixteen_reference_smm_runs(golden_dir, pkg):
    """k_smm (algo='SMM') against 16 outputs of the unmodified reference run with algo='SMM' at order 60
    (tests/golden/rcwa_smm_golden.npz, oracle/make_smm_golden.py), per-structure wavelength / angle / thickness rows."""
    g = np.load(os.path.join(golden_dir, "rcwa_smm_golden.npz"))
    be = pkg.RCWABatch(256, int(g["order"]), max_batch=16, algo='SMM')
    eff, r3, t3, st = be.evaluate_host(g["patterns"], g["wavelength"], g["angle"], g["thickness"], return_orders=True)
    assert not st.any()
    assert np.abs(eff - g["ref_eff"]).max() <= TOL and np.abs(eff - g["ref_eff"]).max() <= TIGHT
    assert np.abs(t3[:, ::-1] - g["ref_tran_cut"]).max() <= TIGHT and np.abs(r3[:, ::-1] - g["ref_refl_cut"]).max() <= TIGHT
    be.close()
