"""GPU (B200): ``pirl_rcwa_eval_static``, the entry point for callers that must not allocate device memory.

Some users of this library cannot call cudaMalloc when they need an efficiency: a rollout worker whose GPU has already been
handed, almost entirely, to the learner framework's memory pool, or a loop that must not trigger the device-wide
synchronisation an allocation implies.  For them the library keeps a fixed workspace inside itself (a static device array,
at most 32 structures of the order-40 DeflectorGym setting per call) and this entry point uses it instead of a handle's
cudaMalloc'ed buffers.

The test calls it directly through the C ABI (ctypes, no handle) and compares with ``pirl_rcwa_eval`` on a handle, twice in a
row with different inputs: the second call runs on a workspace that still holds the first call's matrices, which is exactly
the situation the entry point has to clear up after itself.  It also checks that sizes beyond the stated maximum are refused
with a message rather than overrunning the array.

This file is named ``test_zz_*`` so that pytest collects it last: with ``-x`` a failure here cannot hide any other test.  It
depends on no state of the tests before it (it passes alone in a fresh process as well as after the full suite).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-6        # north_star tolerance, as in tests/test_gpu_parity.py
TIGHT = 1e-9      # regression guard used by the other GPU tests (same kernels on both sides: observed difference is 0)
MAX_BATCH = 32    # PIRL_STATIC_MAX_BATCH in include/pirl_rcwa.h


def _cfg(lib_mod, n_cells, order=40, device=0):
    return lib_mod.PirlCfg(n_cells=n_cells, order=order, n_inc=1.45, n_exit=1.0, n_groove=1.0, theta_deg=0.0, pol=1,
                           max_batch=1, device=device, algo=0, precision=0)


def _static_eval(torch, lib_mod, lib, cfg, pats, rows, stride):
    """One call of the handle-free entry point on torch-owned device buffers (inputs and outputs are the caller's)."""
    B = pats.shape[0]
    dev = pats.device
    eff = torch.full((B,), float("nan"), dtype=torch.float64, device=dev)
    r3 = torch.full((B, 3), float("nan"), dtype=torch.float64, device=dev)
    t3 = torch.full((B, 3), float("nan"), dtype=torch.float64, device=dev)
    st = torch.full((B,), -1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    lib_mod.check(lib.pirl_rcwa_eval_static(C.byref(cfg), pats.data_ptr(), rows.data_ptr(), stride, B, eff.data_ptr(),
                                           r3.data_ptr(), t3.data_ptr(), st.data_ptr(), C.c_void_p(stream)))
    torch.cuda.synchronize()
    return eff.cpu().numpy(), r3.cpu().numpy(), t3.cpu().numpy(), st.cpu().numpy()


def test_static_workspace_entry_matches_handle_entry_twice_in_a_row():
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    import physics_informed_reinforcement_learning_b200 as pkg
    from physics_informed_reinforcement_learning_b200 import _lib as lib_mod
    from physics_informed_reinforcement_learning_b200.materials import n_silicon_real
    lib = lib_mod.load()
    n_cells = 256
    cfg = _cfg(lib_mod, n_cells)
    be = pkg.RCWABatch(n_cells, 40, max_batch=MAX_BATCH)      # the existing entry point, for comparison only
    rng = np.random.default_rng(20261106)
    # call 1: a full workspace (32 structures), per-structure parameter rows; call 2: fewer structures, broadcast row,
    # different patterns -- every buffer of the static array still holds call 1's data when it starts
    cases = []
    B1 = MAX_BATCH
    wl1 = rng.uniform(900.0, 1300.0, B1)
    rows1 = np.stack([wl1, rng.uniform(50.0, 70.0, B1), rng.uniform(250.0, 400.0, B1), n_silicon_real(wl1)], axis=1)
    cases.append((rng.choice(np.array([-1, 1], dtype=np.int8), size=(B1, n_cells)), rows1, 4))
    B2 = 11
    rows2 = np.array([[1100.0, 60.0, 325.0, float(n_silicon_real(np.array([1100.0]))[0])]])
    blocky = np.repeat(rng.choice(np.array([-1, 1], dtype=np.int8), size=(B2, n_cells // 16)), 16, axis=1)
    cases.append((blocky, rows2, 0))
    for pats_np, rows_np, stride in cases:
        pats = torch.from_numpy(np.ascontiguousarray(pats_np)).cuda()
        rows = torch.from_numpy(np.ascontiguousarray(rows_np)).cuda()
        eff_s, r3_s, t3_s, st_s = _static_eval(torch, lib_mod, lib, cfg, pats, rows, stride)
        eff_h, r3_h, t3_h, st_h = be.evaluate(pats, params=rows, return_orders=True)
        torch.cuda.synchronize()
        eff_h, r3_h, t3_h, st_h = eff_h.cpu().numpy(), r3_h.cpu().numpy(), t3_h.cpu().numpy(), st_h.cpu().numpy()
        assert (st_s == 0).all() and (st_h == 0).all(), (st_s, st_h)
        assert np.isfinite(eff_s).all() and np.isfinite(r3_s).all() and np.isfinite(t3_s).all()
        worst = max(np.abs(eff_s - eff_h).max(), np.abs(r3_s - r3_h).max(), np.abs(t3_s - t3_h).max())
        assert worst <= TOL, worst
        assert worst <= TIGHT, worst
        # energy is conserved by a lossless grating: an output that did not come from a real solve would not satisfy this
        assert eff_s.min() >= -1e-12 and eff_s.max() <= 1.0 + 1e-12
    be.close()


def test_static_workspace_entry_rejects_sizes_beyond_its_maximum():
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    from physics_informed_reinforcement_learning_b200 import _lib as lib_mod
    lib = lib_mod.load()
    pats = torch.ones(MAX_BATCH + 1, 256, dtype=torch.int8, device="cuda")
    rows = torch.tensor([[1100.0, 60.0, 325.0, 3.5]], dtype=torch.float64, device="cuda")
    eff = torch.zeros(MAX_BATCH + 1, dtype=torch.float64, device="cuda")

    def call(cfg, B):
        return lib.pirl_rcwa_eval_static(C.byref(cfg), pats.data_ptr(), rows.data_ptr(), 0, B, eff.data_ptr(), None, None,
                                         None, C.c_void_p(torch.cuda.current_stream().cuda_stream))

    assert call(_cfg(lib_mod, 256), MAX_BATCH + 1) == -13            # too many structures for the array
    assert b"PIRL_STATIC_MAX_BATCH" in lib.pirl_last_error()
    assert call(_cfg(lib_mod, 2048), 4) == -13                       # pattern longer than the table region
    assert b"PIRL_STATIC_MAX_CELLS" in lib.pirl_last_error()
    assert call(_cfg(lib_mod, 256, order=60), 4) == -13              # other orders need a handle
    assert call(_cfg(lib_mod, 256), 0) == 0                          # empty batch: nothing launched, nothing touched
    torch.cuda.synchronize()
    assert float(eff.abs().sum()) == 0.0
