"""DeflectorGym-compatible envs on the B200 RCWA backend.

  * ``MeentIndex``          same API as deflector_gym/envs/meent_env.py:66-99 (reset / step / get_efficiency,
                            attrs eff, struct), one env, host arrays -- the drop-in under the existing Q-loop.
  * ``VectorDeflectorEnv``  B envs stepped together: patterns live on the GPU as int8 [B, n]; one batched RCWA
                            launch per step; flip / reward / best tracking are device kernels (bit-exact integer
                            and float64 logic, base.py:37-42, meent_env.py:90-99, wrappers.py:4-23); TimeLimit
                            (main.py:139) and auto-reset included.  ``random_bunch_init`` stays on the host NumPy
                            global RNG so resets consume random numbers exactly like the reference.

``gym`` is optional: when absent (as in this image) light-weight Box/Discrete stand-ins expose the same fields.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib
from .solver import RCWABatch, shared_backend

try:  # pragma: no cover - gym is not installed in the build image
    import gym as _gym
    _Box, _Discrete, _EnvBase = _gym.spaces.Box, _gym.spaces.Discrete, _gym.Env
except Exception:  # noqa: BLE001
    _gym = None

    class _Box:
        def __init__(self, low, high, shape, dtype):
            self.low = np.full(shape, low, dtype=dtype)
            self.high = np.full(shape, high, dtype=dtype)
            self.shape, self.dtype = tuple(shape), np.dtype(dtype)

        def sample(self):
            return np.random.uniform(self.low, self.high).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    class _Discrete:
        def __init__(self, n):
            self.n = int(n)
            self.shape, self.dtype = (), np.dtype(np.int64)

        def sample(self):
            return int(np.random.randint(self.n))

        def contains(self, x):
            return 0 <= int(x) < self.n

    class _EnvBase:
        pass


def random_bunch_init(n_cells=256, mfs=30, mu=0) -> np.ndarray:
    """Host-side initial structure (deflector_gym/envs/utils.py:29-44): all ridge (+1) with isolated grooves;
    consumes exactly one ``np.random.rand()`` per groove from the GLOBAL NumPy RNG, like the reference."""
    sigma = n_cells / 4
    cells = np.ones(n_cells, dtype=np.float64)
    pos = mfs + 1
    while pos < n_cells:
        nxt = pos + mfs + int(abs(mu + np.random.rand() * sigma))
        cells[pos] = -1.0
        pos = nxt + 1
    return cells


class MeentIndex(_EnvBase):
    """One deflector env; ``step(action)`` flips pixel ``action`` and rewards the efficiency change."""

    def __init__(self, n_cells=256, wavelength=1100, desired_angle=70, order=40, thickness=325, alpha=0.05, mfs=1):
        self.n_cells, self.wavelength, self.desired_angle = n_cells, wavelength, desired_angle
        self.order, self.thickness = order, thickness
        self.alpha, self.mfs = 0.05, mfs
        if mfs != 1:
            raise ValueError("mfs > 1 (minimum-feature-size penalty) is dead code in the reference at the pinned "
                             "commit (meent_env.py:48-64 does not compile); only mfs=1 is supported")
        self.struct = None
        self.eff = None
        self.observation_space = _Box(low=-1., high=1., shape=(n_cells,), dtype=np.float64)
        self.action_space = _Discrete(n_cells)
        self._backend = None

    def _be(self) -> RCWABatch:
        if self._backend is None:
            self._backend = shared_backend(self.n_cells, self.order, max_batch=64)
        return self._backend

    def get_efficiency(self, struct) -> float:
        struct = np.asarray(struct, dtype=np.float64)
        eff = self._be().evaluate_host(struct[None, :], float(self.wavelength), float(self.desired_angle),
                                       float(self.thickness))
        return float(eff[0])

    def initialize_struct(self, *args, **kwargs):
        return random_bunch_init(*args, **kwargs)

    def flip(self, pos):
        if 0 <= pos <= (self.n_cells - 1):
            self.struct[pos] = 1 if self.struct[pos] == -1 else -1

    def reset(self):
        # the reference calls initialize_struct() without arguments -> always 256 cells (meent_env.py:85)
        self.struct = self.initialize_struct()
        self.eff = self.get_efficiency(self.struct)
        return self.struct.copy()

    def step(self, action):
        prev_eff = self.eff
        self.flip(action)
        self.eff = self.get_efficiency(self.struct)
        reward = self.eff - prev_eff
        return self.struct.copy(), reward, False, {}


def make(env_id='MeentIndex-v0', **kwargs):
    """``deflector_gym.make`` stand-in for the one env id main.py uses (main.py:130-141)."""
    if env_id != 'MeentIndex-v0':
        raise ValueError(f"{env_id}: only 'MeentIndex-v0' is on the accelerated path")
    return MeentIndex(**kwargs)


_STATUS_BITS = (1, 2, 4)      # PIRL_ST_EIG_NOT_CONVERGED, PIRL_ST_NONFINITE, PIRL_ST_NOT_POSDEF (include/pirl_rcwa.h)


def status_flags(status, mask, acc):
    """Bitwise OR of the per-structure status words, kept as one 0/1 flag per bit of ``mask`` and folded into ``acc``
    (max of 0/1 flags = OR, which also all-reduces across ranks with MAX; the numeric max of the words is NOT their OR)."""
    seen = ((status.reshape(-1, 1) & mask) != 0).any(dim=0).to(acc.dtype)
    return acc.maximum(seen)


class VectorDeflectorEnv:
    """B independent MeentIndex envs advanced by one batched RCWA launch per step (device resident)."""

    def __init__(self, num_envs, n_cells=256, wavelength=1100, desired_angle=60, order=40, thickness=325,
                 max_episode_steps=128, device=None, backend: RCWABatch | None = None, data_dir=None,
                 check_status_every=0):
        import torch
        self.torch = torch
        self.B, self.n_cells = int(num_envs), int(n_cells)
        self.wavelength, self.desired_angle, self.thickness, self.order = wavelength, desired_angle, thickness, order
        self.max_episode_steps = max_episode_steps
        self.backend = backend or RCWABatch(n_cells, order, max_batch=max(64, min(self.B, 8192)), device=device)
        self.device = self.backend.device
        self.lib = _lib.load()
        dev = self.device
        self.patterns = torch.ones(self.B, self.n_cells, dtype=torch.int8, device=dev)
        self.eff = torch.zeros(self.B, dtype=torch.float64, device=dev)
        self.reward = torch.zeros(self.B, dtype=torch.float64, device=dev)
        self.best_eff = torch.zeros(self.B, dtype=torch.float64, device=dev)
        self.best_patterns = torch.ones(self.B, self.n_cells, dtype=torch.int8, device=dev)
        self.elapsed = np.zeros(self.B, dtype=np.int64)
        # best structure over ALL episodes of all envs (BestRecorder keeps it per episode, wrappers.py:10-16, and the
        # reference's callback writes it at every episode end, main.py:76-85): survives the auto-reset
        self.global_best_eff = torch.full((), -1.0, dtype=torch.float64, device=dev)
        self.global_best_pattern = torch.ones(self.n_cells, dtype=torch.int8, device=dev)
        self.episodes_done = 0
        # StructureWriter (utils.py:36-82): the structure is written at every reset as <eff*100>_<timestamp>.npy
        self.data_dir = data_dir
        if data_dir is not None:
            os.makedirs(os.path.join(data_dir, "MeentIndex"), exist_ok=True)
        # solver status words (non-converged eigensolve, non-finite result, not positive definite): checked every N steps
        # (0 = never on the step path; check_status() on demand).  The reference raises LinAlgError instead.
        self.check_status_every = int(check_status_every)
        self._status_acc = torch.zeros(len(_STATUS_BITS), dtype=torch.int32, device=dev)   # one 0/1 flag per status bit
        self._status_mask = torch.tensor(_STATUS_BITS, dtype=torch.int32, device=dev)
        self._steps = 0
        rows, _ = self.backend._params(1, float(wavelength), float(desired_angle), float(thickness), None)
        self.params = torch.from_numpy(rows).to(dev)
        self.observation_space = _Box(low=-1., high=1., shape=(self.n_cells,), dtype=np.float64)
        self.action_space = _Discrete(self.n_cells)

    # -- helpers -------------------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def _evaluate(self, patterns):
        eff, st = self.backend.evaluate(patterns, params=self.params, return_status=True)
        if st.numel():
            self._status_acc = status_flags(st, self._status_mask, self._status_acc)       # device side, no sync
        return eff

    def check_status(self):
        """Raise if any evaluation since the last check set a solver status bit (one device->host read).  With an initialised
        process group of more than one rank this is a COLLECTIVE (the flags are all-reduced first), so that every rank raises
        together instead of one rank leaving the others blocked in the next gradient all_reduce: call it on all ranks."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self._status_acc, op=dist.ReduceOp.MAX)
        bits = int((self._status_acc * self._status_mask).sum().item())
        self._status_acc.zero_()
        if bits:
            raise FloatingPointError(f"RCWA solver status bits {bits:#x} set (1 = eigensolver not converged, 2 = non-finite "
                                     f"efficiency, 4 = permittivity matrices not positive definite)")

    def _fold_global_best(self, idx=None):
        """Fold the per-episode bests (all envs, or the envs ``idx`` that are about to be reset) into the global best."""
        torch = self.torch
        be = self.best_eff if idx is None else self.best_eff[idx]
        bp = self.best_patterns if idx is None else self.best_patterns[idx]
        m, j = be.max(dim=0)
        better = m > self.global_best_eff
        self.global_best_eff = torch.where(better, m, self.global_best_eff)
        self.global_best_pattern = torch.where(better, bp[j], self.global_best_pattern)

    def best(self):
        """(efficiency, structure float64 [n]) of the best structure seen so far, finished and running episodes alike."""
        self._fold_global_best()
        return float(self.global_best_eff.item()), self.global_best_pattern.to(self.torch.float64).cpu().numpy()

    def _write_structures(self, idx, eff, patterns):
        from datetime import datetime
        stamp = datetime.now().strftime('%Y%m%d-%H%M%S')
        e, p = eff.cpu().numpy(), patterns.cpu().numpy().astype(np.float64)
        for q, i in enumerate(idx):
            name = f'{e[q] * 100:.6f}'.replace('.', '-') + '_' + stamp + f'_e{int(i)}'
            np.save(os.path.join(self.data_dir, "MeentIndex", name), p[q])

    def _host_init(self, count):
        if self.n_cells != 256:
            # MeentIndex.reset ignores n_cells (always the 256-cell default); other sizes use the same rule scaled
            return np.stack([random_bunch_init(self.n_cells) for _ in range(count)]).astype(np.int8)
        return np.stack([random_bunch_init() for _ in range(count)]).astype(np.int8)

    def obs(self):
        """float64 [B, n] observation like ``struct.copy()`` (device tensor)."""
        return self.patterns.to(self.torch.float64)

    # -- gym-like API --------------------------------------------------------------------------------
    def reset(self, indices=None):
        torch = self.torch
        if indices is None:
            idx = np.arange(self.B)
        else:
            idx = np.asarray(indices, dtype=np.int64)
        if idx.size == 0:
            return self.obs()
        init = torch.from_numpy(self._host_init(idx.size)).to(self.device)
        if idx.size == self.B:
            self.patterns.copy_(init)
            eff = self._evaluate(self.patterns)
            self.eff.copy_(eff)
            self.best_eff.copy_(eff)
            self.best_patterns.copy_(self.patterns)
        else:
            tidx = torch.from_numpy(idx).to(self.device)
            self.patterns[tidx] = init
            eff = self._evaluate(init)
            self.eff[tidx] = eff
            self.best_eff[tidx] = eff
            self.best_patterns[tidx] = init
        self.elapsed[idx] = 0
        if self.data_dir is not None:
            self._write_structures(idx, eff, init)
        return self.obs()

    def step(self, actions, auto_reset=True):
        """actions: int64 [B] (tensor or array).  Returns (obs[B,n] f64, reward[B] f64, done[B] bool, info)."""
        torch = self.torch
        if not torch.is_tensor(actions):
            actions = torch.as_tensor(np.asarray(actions, dtype=np.int64))
        actions = actions.to(self.device, dtype=torch.int64).contiguous()
        st = self._stream()
        _lib.check(self.lib.pirl_env_flip(self.patterns.data_ptr(), actions.data_ptr(), self.B, self.n_cells, st))
        eff_new = self._evaluate(self.patterns)
        _lib.check(self.lib.pirl_env_reward(eff_new.data_ptr(), self.eff.data_ptr(), self.reward.data_ptr(),
                                            self.best_eff.data_ptr(), self.best_patterns.data_ptr(),
                                            self.patterns.data_ptr(), self.B, self.n_cells, st))
        self.elapsed += 1
        done_np = np.zeros(self.B, dtype=bool)
        if self.max_episode_steps is not None:
            done_np = self.elapsed >= self.max_episode_steps
        reward = self.reward.clone()
        obs = self.obs()
        info = {"TimeLimit.truncated": done_np.copy()}
        self._steps += 1
        if self.check_status_every and self._steps % self.check_status_every == 0:
            self.check_status()
        if auto_reset and done_np.any():
            didx = np.nonzero(done_np)[0]
            tidx = torch.from_numpy(didx).to(self.device)
            info["terminal_eff"] = self.eff.clone()
            # episode results of the envs that finish now (BestRecorder's env.best at episode end, main.py:76-85)
            info["episode_env"] = didx
            info["episode_best_eff"] = self.best_eff[tidx].clone()
            info["episode_best_patterns"] = self.best_patterns[tidx].clone()
            self._fold_global_best(tidx)
            self.episodes_done += int(didx.size)
            self.reset(didx)
            obs = self.obs()
        return obs, reward, torch.from_numpy(done_np).to(self.device), info

    def get_efficiency_batch(self, structs):
        """structs: [B, n] of +-1 (numpy or tensor) -> float64 [B] on the device."""
        torch = self.torch
        if not torch.is_tensor(structs):
            structs = torch.as_tensor(np.asarray(structs))
        pat = torch.where(structs.to(self.device) > 0, 1, -1).to(torch.int8)
        return self._evaluate(pat)
