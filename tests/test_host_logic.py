"""CPU: host-side logic -- env transitions vs the reference's own random_bunch_init draws, shard arithmetic,
parameter packing, ABI surface (no GPU compute)."""
import ctypes
import os
import re

import numpy as np
import pytest

from oracle import env_oracle
from physics_informed_reinforcement_learning_b200 import _lib, envs, materials
from physics_informed_reinforcement_learning_b200.build import LIB_PATH, build_library
from physics_informed_reinforcement_learning_b200.dist import shard_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bunch_init_matches_reference_draws(golden_dir):
    g = np.load(os.path.join(golden_dir, "env_golden.npz"))
    for seed, ref in zip(g["seeds"], g["ref_bunch"]):
        np.random.seed(int(seed))
        a = envs.random_bunch_init()
        np.random.seed(int(seed))
        b = env_oracle.bunch_init()
        assert np.array_equal(a.astype(np.int8), ref)
        assert np.array_equal(b.astype(np.int8), ref)
        assert a.dtype == np.float64
    np.random.seed(42)
    assert list(np.nonzero(envs.random_bunch_init() == -1)[0]) == [31, 85, 176, 253]


def test_flip_rule():
    s = np.ones(8)
    for pos, expect in [(0, -1), (0, 1), (7, -1), (8, None), (-1, None)]:
        before = s.copy()
        env_oracle.flip(s, pos)
        if expect is None:
            assert np.array_equal(s, before)
        else:
            assert s[pos] == expect


def test_scalar_env_semantics_with_stub_efficiency():
    calls = []

    def eff_fn(struct):
        calls.append(struct.copy())
        return float((struct == -1).sum()) / 256.0

    np.random.seed(3)
    env = env_oracle.ScalarEnv(eff_fn, max_episode_steps=3)
    obs = env.reset()
    e0 = env.eff
    obs, r, done, info = env.step(0)
    assert r == env.eff - e0 and not done and obs[0] == -1
    obs, r, done, info = env.step(0)
    assert obs[0] == 1 and env.best[0] >= env.eff
    obs, r, done, info = env.step(300)      # out of range: no-op, still one evaluation
    assert r == 0.0 and done and info["TimeLimit.truncated"]
    assert len(calls) == 4


def test_shard_range_covers_everything():
    for total in (0, 1, 7, 4096, 16384, 1000):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_material_lookup_contract():
    assert float(materials.find_nk_index("p_si__real", 1100)) == pytest.approx(3.551726470588235, abs=1e-14)
    nk = materials.find_nk_index("p_si", 500.0)
    assert np.iscomplexobj(nk) and nk.imag > 0
    assert materials.find_nk_index(1.45, 1100) == 1.45
    with pytest.raises(KeyError):
        materials.find_nk_index("unobtainium", 1100)


def test_library_exports_every_declared_symbol():
    build_library()
    assert os.path.exists(LIB_PATH)
    header = open(os.path.join(ROOT, "include", "pirl_rcwa.h")).read()
    declared = set(re.findall(r"\b(pirl_[a-z_0-9]+)\s*\(", header))
    declared -= {"pirl_handle", "pirl_cfg"}
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    lib = ctypes.CDLL(LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert _lib.load().pirl_abi_version() == 2


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    cfg = _lib.PirlCfg(n_cells=256, order=40, n_inc=1.45, n_exit=1.0, n_groove=1.0, theta_deg=0.0, pol=1,
                       max_batch=8, device=0)
    h = ctypes.c_void_p()
    rc = lib.pirl_rcwa_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc != 0 and b"CPU fallback" in lib.pirl_last_error()
    from physics_informed_reinforcement_learning_b200 import RCWABatch
    with pytest.raises(RuntimeError):
        RCWABatch(256, 40)


def test_argument_errors_reported_before_touching_the_device():
    lib = _lib.load()
    h = ctypes.c_void_p()
    for kw in (dict(pol=2), dict(theta_deg=10.0), dict(order=0), dict(order=200), dict(max_batch=0), dict(algo=2),
               dict(precision=1)):
        base = dict(n_cells=256, order=40, n_inc=1.45, n_exit=1.0, n_groove=1.0, theta_deg=0.0, pol=1, max_batch=8,
                    device=0)
        base.update(kw)
        cfg = _lib.PirlCfg(**base)
        assert lib.pirl_rcwa_create(ctypes.byref(cfg), ctypes.byref(h)) < 0, kw
    assert lib.pirl_rcwa_create(None, ctypes.byref(h)) < 0


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "physics_informed_reinforcement_learning_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("oracle/make_golden.py", ""), fn


def test_status_words_are_ored_bit_by_bit():
    """A batch holding statuses 1 and 4 must report 5 (the numeric max, 4, would drop the eigensolver bit)."""
    import torch
    from physics_informed_reinforcement_learning_b200.envs import status_flags, _STATUS_BITS
    mask = torch.tensor(_STATUS_BITS, dtype=torch.int32)
    acc = torch.zeros(3, dtype=torch.int32)
    acc = status_flags(torch.tensor([0, 1, 0, 4], dtype=torch.int32), mask, acc)
    assert int((acc * mask).sum()) == 5
    acc = status_flags(torch.tensor([0, 0], dtype=torch.int32), mask, acc)          # flags are sticky until cleared
    assert int((acc * mask).sum()) == 5
    acc = status_flags(torch.tensor([2 | 0x100], dtype=torch.int32), mask, acc)      # internal high bits are ignored
    assert int((acc * mask).sum()) == 7


def test_env_oracle_reproduces_reference_step_fixtures(golden_dir):
    """oracle/env_oracle.py against trajectories produced by the reference's own flip / reset / step method objects
    (oracle/make_env_step_golden.py): states, efficiencies and rewards bit for bit, out-of-range actions included."""
    from oracle.make_env_step_golden import stub_efficiency
    g = np.load(os.path.join(golden_dir, "env_step_golden.npz"))
    for k, seed in enumerate(g["seeds"]):
        np.random.seed(int(seed))
        env = env_oracle.ScalarEnv(stub_efficiency)
        assert np.array_equal(env.reset(), g["init"][k]) and env.eff == g["effs"][k][0]
        for t, a in enumerate(g["actions"][k]):
            s, r, d, _ = env.step(int(a))
            assert np.array_equal(s, g["states"][k][t])
            assert r == g["rewards"][k][t] and env.eff == g["effs"][k][t + 1] and d == bool(g["dones"][k][t])
