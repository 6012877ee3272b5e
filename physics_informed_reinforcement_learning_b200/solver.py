"""Host-side mirror of the reference's solver interface, backed by the C-ABI CUDA library.

  * ``RCWABatch``           one handle (device workspaces) per (device, n_cells, order); batched evaluation of
                            +-1 patterns on device tensors (``evaluate``) or host arrays (``evaluate_host``).
  * ``JLABCode``            same constructor / ``reproduce_acs_cell`` contract as the reference adapter
                            (deflector-gym/third_party/meent/JLAB/solver.py:8-38) for one structure.

Nothing here computes RCWA on the CPU: every number comes from libpirl_rcwa.so (sm_100a kernels).
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import _lib
from .materials import find_nk_index, n_silicon_real

ST_EIG_NOT_CONVERGED = 1
ST_NONFINITE = 2
ST_NOT_POSDEF = 4


def _torch():
    import torch
    return torch


class RCWABatch:
    """Batched 1D TM RCWA of binary deflectors on one B200 (replaces JLABCode + RCWALight for B structures)."""

    def __init__(self, n_cells=256, order=40, n_I=1.45, n_II=1.0, n_groove=1.0, theta=0.0, pol=1,
                 max_batch=4096, device=None, algo='TMM'):
        torch = _torch()
        if not torch.cuda.is_available():
            raise RuntimeError("RCWABatch needs a CUDA device (B200); there is no CPU fallback")
        self.lib = _lib.load()
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device("cuda", device if isinstance(device, int) else torch.device(device).index or 0)
        self.n_cells, self.order, self.ff = int(n_cells), int(order), 2 * int(order) + 1
        self.n_I, self.n_II, self.n_groove = float(n_I), float(n_II), float(n_groove)
        self.pol = int(pol)
        if algo not in ('TMM', 'SMM', 0, 1):
            raise ValueError(algo)                       # rcwa.py:31-32
        self.algo = 1 if algo in ('SMM', 1) else 0       # 'TMM': enhanced transmittance matrices; 'SMM': Redheffer star products
        cfg = _lib.PirlCfg(n_cells=self.n_cells, order=self.order, n_inc=self.n_I, n_exit=self.n_II,
                           n_groove=self.n_groove, theta_deg=float(theta), pol=int(pol), max_batch=int(max_batch),
                           device=self.device.index, algo=self.algo, precision=0)
        h = C.c_void_p()
        _lib.check(self.lib.pirl_rcwa_create(C.byref(cfg), C.byref(h)))
        self._h = h
        self.max_batch = int(max_batch)

    def close(self):
        if getattr(self, "_h", None):
            self.lib.pirl_rcwa_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------------
    def _params(self, B, wavelength, angle, thickness, n_ridge):
        """[1,4] (broadcast) or [B,4] float64 rows {wavelength, angle, thickness, n_ridge}; n_ridge defaults to
        'p_si__real' interpolated on the host (convolution_matrix.py:40-52)."""
        wl = np.atleast_1d(np.asarray(wavelength, dtype=np.float64))
        if n_ridge is None:
            n_ridge = n_silicon_real(wl)
        cols = [wl, np.atleast_1d(np.asarray(angle, dtype=np.float64)),
                np.atleast_1d(np.asarray(thickness, dtype=np.float64)),
                np.atleast_1d(np.asarray(n_ridge, dtype=np.float64))]
        n = max(c.size for c in cols)
        if n not in (1, B):
            raise ValueError(f"per-structure parameters must have length 1 or B={B}, got {n}")
        rows = np.empty((n, 4), dtype=np.float64)
        for j, c in enumerate(cols):
            if c.size not in (1, n):
                raise ValueError("parameter arrays must share one length")
            rows[:, j] = c
        return rows, (4 if n > 1 else 0)

    def _check_params(self, params, B):
        torch = _torch()
        if (not torch.is_tensor(params) or params.dtype != torch.float64 or params.dim() != 2 or params.shape[1] != 4
                or params.shape[0] not in (1, B) or not params.is_cuda):
            raise ValueError(f"params must be a float64 CUDA tensor of shape [1, 4] or [{B}, 4] "
                             "(rows {wavelength, angle, thickness, n_ridge})")
        return params.contiguous()

    def evaluate(self, patterns, wavelength=1100.0, angle=60.0, thickness=325.0, n_ridge=None,
                 return_orders=False, params=None, return_status=False):
        """patterns: torch int8 CUDA tensor [B, n_cells] of +-1.  Returns eff [B] float64 (CUDA), and with
        ``return_orders`` also (de_r3, de_t3, status) -- rows are de[N-1], de[N], de[N+1].
        Asynchronous on the current torch stream; no host<->device copies of bulk data."""
        torch = _torch()
        if patterns.dtype != torch.int8 or not patterns.is_cuda or patterns.dim() != 2:
            raise TypeError("patterns must be a 2-D int8 CUDA tensor")
        if patterns.shape[1] != self.n_cells:
            raise ValueError(f"expected {self.n_cells} cells per pattern")
        patterns = patterns.contiguous()
        B = patterns.shape[0]
        dev = patterns.device
        if params is None:
            rows, stride = self._params(B, wavelength, angle, thickness, n_ridge)
            params = torch.from_numpy(rows).to(dev, non_blocking=True)
        else:
            params = self._check_params(params, B)
            stride = 4 if params.shape[0] > 1 else 0
        eff = torch.empty(B, dtype=torch.float64, device=dev)
        st = torch.empty(B, dtype=torch.int32, device=dev)
        r3 = t3 = None
        if return_orders:
            r3 = torch.empty(B, 3, dtype=torch.float64, device=dev)
            t3 = torch.empty(B, 3, dtype=torch.float64, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(self.lib.pirl_rcwa_eval(self._h, patterns.data_ptr(), params.data_ptr(), stride, B,
                                           eff.data_ptr(), r3.data_ptr() if return_orders else None,
                                           t3.data_ptr() if return_orders else None, st.data_ptr(),
                                           C.c_void_p(stream)))
        self._keep = (patterns, params)  # keep inputs alive until the stream has consumed them
        if return_orders:
            return eff, r3, t3, st
        if return_status:
            return eff, st
        return eff

    def evaluate_host(self, patterns, wavelength=1100.0, angle=60.0, thickness=325.0, n_ridge=None,
                      return_orders=False, out=None, status_out=None):
        """patterns: numpy [B, n_cells] (any real dtype, >0 = ridge).  Host buffers in, host buffers out;
        the H2D/D2H copies happen inside the library call (``pirl_rcwa_eval_host``).  ``out`` / ``status_out``:
        optional preallocated float64 [B] / int32 [B] result buffers (e.g. pinned memory)."""
        pat = np.asarray(patterns)
        if pat.ndim != 2 or pat.shape[1] != self.n_cells:
            raise ValueError(f"patterns must be [B, {self.n_cells}]")
        if pat.dtype != np.int8 or not pat.flags.c_contiguous:
            pat = np.ascontiguousarray(np.where(pat > 0, 1, -1).astype(np.int8))
        B = pat.shape[0]
        rows, stride = self._params(B, wavelength, angle, thickness, n_ridge)
        eff = np.empty(B, dtype=np.float64) if out is None else out
        st = np.empty(B, dtype=np.int32) if status_out is None else status_out
        if eff.dtype != np.float64 or eff.shape != (B,) or st.dtype != np.int32 or st.shape != (B,):
            raise ValueError("out must be float64 [B] and status_out int32 [B]")
        r3 = np.empty((B, 3), dtype=np.float64) if return_orders else None
        t3 = np.empty((B, 3), dtype=np.float64) if return_orders else None
        _lib.check(self.lib.pirl_rcwa_eval_host(
            self._h, pat.ctypes.data, rows.ctypes.data, stride, B, eff.ctypes.data,
            r3.ctypes.data if return_orders else None, t3.ctypes.data if return_orders else None, st.ctypes.data))
        if return_orders:
            return eff, r3, t3, st
        return eff

    def evaluate_stack_host(self, patterns, wavelength=1100.0, angle=60.0, thicknesses=(325.0,), n_ridge=None,
                            return_orders=False):
        """Layer stacks (the reference's layer loop, _base.py:141): patterns numpy [B, L, n_cells] (>0 = ridge), layers in the
        reference's list order; ``thicknesses`` [L] (shared) or [B, L].  Host buffers in and out (``pirl_rcwa_eval_stack_host``).
        The layer ORDER convention follows the handle's ``algo`` like the reference does ('SMM' = reversed list)."""
        pats = np.asarray(patterns)
        if pats.ndim != 3 or pats.shape[2] != self.n_cells:
            raise ValueError(f"patterns must be [B, L, {self.n_cells}]")
        B, L = pats.shape[0], pats.shape[1]
        p8 = np.ascontiguousarray(np.where(pats > 0, 1, -1).astype(np.int8))
        rows, stride = self._params(B, wavelength, angle, 0.0, n_ridge)
        th = np.ascontiguousarray(np.asarray(thicknesses, dtype=np.float64))
        if th.shape == (L,):
            tstride = 0
        elif th.shape == (B, L):
            tstride = L
        else:
            raise ValueError(f"thicknesses must have shape ({L},) or ({B}, {L})")
        eff = np.empty(B, dtype=np.float64)
        st = np.empty(B, dtype=np.int32)
        r3 = np.empty((B, 3), dtype=np.float64) if return_orders else None
        t3 = np.empty((B, 3), dtype=np.float64) if return_orders else None
        _lib.check(self.lib.pirl_rcwa_eval_stack_host(
            self._h, p8.ctypes.data, rows.ctypes.data, stride, th.ctypes.data, tstride, L, B, eff.ctypes.data,
            r3.ctypes.data if return_orders else None, t3.ctypes.data if return_orders else None, st.ctypes.data))
        if return_orders:
            return eff, r3, t3, st
        return eff

    def evaluate_grad(self, patterns, wavelength=1100.0, angle=60.0, thickness=325.0, n_ridge=None, params=None):
        """Efficiency and its adjoint gradient d eff / d s_p.  patterns: torch int8 CUDA tensor [B, n_cells] of +-1.
        Returns (eff [B], grad [B, n_cells], status [B]) as CUDA tensors, asynchronous on the current torch stream.
        The reference has no gradient; the contract is the finite difference of its forward solver w.r.t. one continuous
        cell value (solver.py:21-22), see ``pirl_rcwa_eval_grad`` in include/pirl_rcwa.h."""
        torch = _torch()
        if patterns.dtype != torch.int8 or not patterns.is_cuda or patterns.dim() != 2:
            raise TypeError("patterns must be a 2-D int8 CUDA tensor")
        if patterns.shape[1] != self.n_cells:
            raise ValueError(f"expected {self.n_cells} cells per pattern")
        patterns = patterns.contiguous()
        B = patterns.shape[0]
        dev = patterns.device
        if params is None:
            rows, stride = self._params(B, wavelength, angle, thickness, n_ridge)
            params = torch.from_numpy(rows).to(dev, non_blocking=True)
        else:
            params = self._check_params(params, B)
            stride = 4 if params.shape[0] > 1 else 0
        eff = torch.empty(B, dtype=torch.float64, device=dev)
        grad = torch.empty(B, self.n_cells, dtype=torch.float64, device=dev)
        st = torch.empty(B, dtype=torch.int32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(self.lib.pirl_rcwa_eval_grad(self._h, patterns.data_ptr(), params.data_ptr(), stride, B,
                                                eff.data_ptr(), grad.data_ptr(), st.data_ptr(), C.c_void_p(stream)))
        self._keep = (patterns, params)
        return eff, grad, st

    def evaluate_grad_host(self, patterns, wavelength=1100.0, angle=60.0, thickness=325.0, n_ridge=None):
        """Host buffers in / out (``pirl_rcwa_eval_grad_host``): returns (eff [B], grad [B, n_cells], status [B]) as numpy."""
        pat = np.asarray(patterns)
        if pat.ndim != 2 or pat.shape[1] != self.n_cells:
            raise ValueError(f"patterns must be [B, {self.n_cells}]")
        if pat.dtype != np.int8 or not pat.flags.c_contiguous:
            pat = np.ascontiguousarray(np.where(pat > 0, 1, -1).astype(np.int8))
        B = pat.shape[0]
        rows, stride = self._params(B, wavelength, angle, thickness, n_ridge)
        eff = np.empty(B, dtype=np.float64)
        grad = np.empty((B, self.n_cells), dtype=np.float64)
        st = np.empty(B, dtype=np.int32)
        _lib.check(self.lib.pirl_rcwa_eval_grad_host(self._h, pat.ctypes.data, rows.ctypes.data, stride, B,
                                                     eff.ctypes.data, grad.ctypes.data, st.ctypes.data))
        return eff, grad, st

    def launch_count(self) -> int:
        return int(self.lib.pirl_rcwa_launch_count(self._h))

    def set_stage_timing(self, enable: bool):
        """Bracket every kernel of ``evaluate`` with CUDA events on the launching stream (bench.py's roofline)."""
        _lib.check(self.lib.pirl_rcwa_set_stage_timing(self._h, 1 if enable else 0))

    def set_subchunk(self, structures_per_subchunk: int):
        """-1 automatic, 0 off, n > 0 structures per sub-chunk of the internal pipeline (``pirl_rcwa_set_subchunk``)."""
        _lib.check(self.lib.pirl_rcwa_set_subchunk(self._h, int(structures_per_subchunk)))

    def stage_times(self):
        """{kernel name: (total ms, launches)} since the last call; synchronises the recorded events."""
        n = self.lib.pirl_rcwa_num_stages()
        ms = np.zeros(n, dtype=np.float64)
        cnt = np.zeros(n, dtype=np.int64)
        _lib.check(self.lib.pirl_rcwa_stage_times(self._h, ms.ctypes.data, cnt.ctypes.data))
        return {self.lib.pirl_rcwa_stage_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n)}

    def debug_fetch(self, what: str, B: int):
        ff = self.ff
        code, shape = {"C": (0, (B, ff, ff)), "tri": (1, (B, 2, ff)), "lam": (2, (B, ff)),
                       "W": (3, (B, ff, ff)), "nrot": (4, (B,)), "wstatus": (7, (B,)), "phase_clk": (5, (32,)), "trace": (6, (0,))}[what]
        if what == "trace":
            raw = np.zeros(1 + 4 * (1 << 20), dtype=np.uint64)
            _lib.check(self.lib.pirl_rcwa_debug_fetch(self._h, 6, raw.ctypes.data, raw.size))
            n = int(min(raw[0], 1 << 20))
            return raw[1:1 + 4 * n].reshape(n, 4)
        out = np.empty(shape, dtype=np.float64)
        _lib.check(self.lib.pirl_rcwa_debug_fetch(self._h, code, out.ctypes.data, out.size))
        return out


# ----------------------------------------------------------------------------------------------------
_backends = {}
_backends_lock = threading.Lock()


def shared_backend(n_cells, order, n_I=1.45, n_II=1.0, n_groove=1.0, max_batch=4096, device=None, algo='TMM', pol=1) -> RCWABatch:
    """One cached handle per configuration (the reference builds a solver object per call; here the
    workspaces are allocated once)."""
    torch = _torch()
    dev = torch.cuda.current_device() if device is None else device
    key = (int(n_cells), int(order), float(n_I), float(n_II), float(n_groove), int(max_batch), dev, algo, int(pol))
    with _backends_lock:
        if key not in _backends:
            _backends[key] = RCWABatch(n_cells, order, n_I, n_II, n_groove, pol=pol, max_batch=max_batch, device=dev, algo=algo)
        return _backends[key]


class JLABCode:
    """Drop-in for the reference adapter (JLAB/solver.py:8-38), one structure per object.

    Same arguments and error behaviour for the 1D TM case DeflectorGym drives; unsupported modes raise
    ``ValueError`` like the reference does for unknown ``grating_type`` / ``pol`` / ``algo``."""

    def __init__(self, grating_type=0, n_I=1.45, n_II=1., theta=0, phi=0, psi=0, fourier_order=40, period=100,
                 wls=np.linspace(900, 900, 1), pol=1, patterns=None, ucell=None, ucell_materials=None,
                 thickness=(325,), algo='TMM'):
        if grating_type != 0:
            raise ValueError("only grating_type=0 (1D) is on the accelerated path")
        if pol not in (0, 1):
            raise ValueError(pol)                        # _base.py:160-161
        self.pol = pol
        if algo not in ('TMM', 'SMM'):
            raise ValueError(algo)
        if theta != 0 or phi != 0:
            raise ValueError("only normal incidence is on the accelerated path")
        self.algo = algo            # 'SMM' runs the scattering-matrix kernel (k_smm), 'TMM' the transmittance-matrix tails
        self.n_I, self.n_II = n_I, n_II
        self.fourier_order = fourier_order
        self.period = np.atleast_1d(np.asarray(period, dtype=np.float64))
        self.wls = np.atleast_1d(np.asarray(wls, dtype=np.float64))
        self.ucell = np.asarray(ucell, dtype=np.float64)
        self.thickness = np.atleast_1d(np.asarray(thickness, dtype=np.float64))
        if self.ucell.ndim != 3 or self.ucell.shape[1] != 1:
            raise ValueError("ucell must have shape (n_layers, 1, n_cells): 1D layers")
        if self.thickness.size != self.ucell.shape[0] or self.wls.size != 1:
            raise ValueError("one thickness per layer / one wavelength per call")

    def reproduce_acs_cell(self, n_ridge, n_groove):
        wl = float(self.wls[0])
        if isinstance(n_ridge, str):
            n_ridge = find_nk_index(n_ridge, wl)
        if np.iscomplexobj(n_ridge):
            raise ValueError("lossy ridge index is not on the accelerated path (use '<material>__real')")
        struct = self.ucell[0, 0]
        # period = wl / sin(angle)  ->  the kernels take the deflection angle
        ratio = wl / float(self.period[0])
        angle = np.degrees(np.arcsin(ratio))
        be = shared_backend(struct.size, self.fourier_order, self.n_I, self.n_II, float(n_groove), max_batch=64, algo=self.algo,
                            pol=self.pol)
        if self.ucell.shape[0] > 1:          # layer stack (_base.py:141), layers in list order
            eff, r3, t3, st = be.evaluate_stack_host(self.ucell[None, :, 0, :], wl, angle, self.thickness, float(n_ridge),
                                                     return_orders=True)
        else:
            eff, r3, t3, st = be.evaluate_host(struct[None, :], wl, angle, float(self.thickness[0]), float(n_ridge),
                                               return_orders=True)
        if st[0] & (ST_EIG_NOT_CONVERGED | ST_NOT_POSDEF):
            raise np.linalg.LinAlgError(f"RCWA solve failed, status {int(st[0])}")
        return float(eff[0]), r3[0, ::-1].copy(), t3[0, ::-1].copy()
