"""All-core timing harness around the CPU oracle  --  TEST / BENCH INFRASTRUCTURE, NOT PRODUCT.

``bench.py`` uses it for the ``cpu_baseline`` object and for ``--impl reference``: one worker process per host core,
one BLAS thread per worker (multi-threaded BLAS buys nothing at ff=81, so the fair all-core CPU
figure is process-parallel), every worker running ``oracle.rcwa_oracle.efficiency`` -- the NumPy restatement of the
reference's own formulation (same LAPACK zgeev / zgesv / pocketfft calls as meent's ``solve_1d``).
"""
from __future__ import annotations

import os
import time

import numpy as np


def host_cores() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:  # pragma: no cover
        return max(1, os.cpu_count() or 1)


def _init_worker():
    os.environ["OPENBLAS_NUM_THREADS"] = "1"
    os.environ["OMP_NUM_THREADS"] = "1"
    os.environ["MKL_NUM_THREADS"] = "1"
    try:
        from threadpoolctl import threadpool_limits
        global _limit
        _limit = threadpool_limits(limits=1)
    except Exception:  # noqa: BLE001
        pass
    from oracle import rcwa_oracle  # noqa: F401  (import cost outside the timed region)


def _eval_chunk(args):
    pats, wavelength, angle, order, thickness = args
    from oracle import rcwa_oracle as O
    return np.array([O.efficiency(p.astype(np.float64), wavelength, angle, order, thickness) for p in pats])


class OraclePool:
    """Persistent pool so that process start-up and imports stay outside every timed region."""

    def __init__(self, cores: int | None = None):
        import multiprocessing as mp
        self.cores = cores or host_cores()
        for k in ("OPENBLAS_NUM_THREADS", "OMP_NUM_THREADS", "MKL_NUM_THREADS"):
            os.environ.setdefault(k, "1")      # inherited by the spawned workers before they import numpy
        self.pool = mp.get_context("spawn").Pool(self.cores, initializer=_init_worker)
        self.pool.map(_noop, range(self.cores * 2))

    def run(self, patterns, wavelength, angle, order, thickness):
        """Evaluate all patterns; returns (efficiencies, wall seconds)."""
        chunks = [c for c in np.array_split(np.asarray(patterns), self.cores * 4) if len(c)]
        t0 = time.perf_counter()
        out = self.pool.map(_eval_chunk, [(c, wavelength, angle, order, thickness) for c in chunks])
        dt = time.perf_counter() - t0
        return np.concatenate(out), dt

    def close(self):
        self.pool.close()
        self.pool.join()


def _noop(_):
    return 0


def single_eval_seconds(pattern, wavelength, angle, order, thickness, reps=3) -> float:
    from oracle import rcwa_oracle as O
    O.efficiency(pattern.astype(np.float64), wavelength, angle, order, thickness)
    t0 = time.perf_counter()
    for _ in range(reps):
        O.efficiency(pattern.astype(np.float64), wavelength, angle, order, thickness)
    return (time.perf_counter() - t0) / reps
