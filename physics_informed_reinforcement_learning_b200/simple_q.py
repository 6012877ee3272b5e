"""Native SimpleQ loop over the vectorised deflector env.

Replaces the un-vendored RLlib driver of the reference (main.py:146-186, configs/simple_q.py:47-82) with a thin PyTorch loop
around the unchanged Q-network hook; the env physics is the batched RCWA backend.  Hyper-parameters are the reference's
``multiple_worker`` config: epsilon-greedy 0.99 -> 0.01 over 100 k timesteps, replay 100 k with learning start at 1000,
train batch 512, Adam lr 1e-3, gamma 0.99, target update every 2000 timesteps with tau 0.1, 128-step episodes
(main.py:139 TimeLimit).  RLlib's exact loss/replay internals are not in the reference tree -> RL-loop parity is unpinned
(SURVEY.md section 8c); what is pinned and tested is the env transition and the hyper-parameter schedule.

Multi-GPU (one process per GPU): envs and replay are rank-local; ranks stay in lock step by averaging gradients
(``all_reduce``) after an initial ``broadcast`` of the policy weights; ``max_efficiency`` is reduced with MAX.  No env data
crosses GPUs.
"""
from __future__ import annotations

import copy
import dataclasses
import os
import time

import numpy as np
import torch
import torch.distributed as dist
from torch import nn


@dataclasses.dataclass
class SimpleQConfig:
    gamma: float = 0.99                        # configs/simple_q.py:67
    lr: float = 1e-3                           # :66
    train_batch_size: int = 512                # :68 (global, split over ranks)
    replay_capacity: int = 100_000             # :60
    learning_starts: int = 1000                # :59
    target_network_update_freq: int = 2000     # :53
    tau: float = 0.1                           # :69
    initial_epsilon: float = 0.99              # :78
    final_epsilon: float = 0.01                # :79
    epsilon_timesteps: int = 100_000           # :80
    timesteps_per_update: int = 32             # :71-73: 16 rollout workers x rollout_fragment_length 2 sampled per train batch
    max_updates_per_iter: int = 64             # cap on learner updates after one vector-env step (B * world timesteps)
    seed: int = 42                             # utils.py:14-18 seed_all
    cuda_graph: bool = True                    # replay forward + backward of the learner step from a CUDA graph (CUDA only)


def epsilon_at(cfg: SimpleQConfig, timestep: int) -> float:
    """Linear schedule of RLlib's EpsilonGreedy (PiecewiseSchedule between (0, initial) and (epsilon_timesteps, final))."""
    f = min(max(timestep / float(cfg.epsilon_timesteps), 0.0), 1.0)
    return cfg.initial_epsilon + f * (cfg.final_epsilon - cfg.initial_epsilon)


class DeviceReplay:
    """Uniform ring buffer resident on the env's device: obs / next_obs as int8 patterns (256 B per transition)."""

    def __init__(self, capacity: int, n_cells: int, device):
        self.capacity, self.size, self.pos = int(capacity), 0, 0
        self.obs = torch.empty(capacity, n_cells, dtype=torch.int8, device=device)
        self.next_obs = torch.empty(capacity, n_cells, dtype=torch.int8, device=device)
        self.action = torch.empty(capacity, dtype=torch.int64, device=device)
        self.reward = torch.empty(capacity, dtype=torch.float32, device=device)
        self.done = torch.empty(capacity, dtype=torch.bool, device=device)

    def add(self, obs, action, reward, next_obs, done):
        n = obs.shape[0]
        idx = (torch.arange(n, device=obs.device) + self.pos) % self.capacity
        self.obs[idx] = obs.to(torch.int8)
        self.next_obs[idx] = next_obs.to(torch.int8)
        self.action[idx] = action
        self.reward[idx] = reward.to(torch.float32)
        self.done[idx] = done
        self.pos = (self.pos + n) % self.capacity
        self.size = min(self.size + n, self.capacity)

    def sample(self, batch: int, generator=None):
        idx = torch.randint(0, self.size, (batch,), device=self.obs.device, generator=generator)
        return self.obs[idx], self.action[idx], self.reward[idx], self.next_obs[idx], self.done[idx]


class SimpleQTrainer:
    """env: object with ``B``, ``n_cells``, ``device``, ``reset() -> obs[B,n]``, ``step(actions) -> (obs, reward, done, info)``
    and ``best_eff`` (``VectorDeflectorEnv`` or a test double)."""

    def __init__(self, env, model: nn.Module, cfg: SimpleQConfig | None = None, log_dir: str | None = None):
        self.env, self.cfg = env, cfg or SimpleQConfig()
        self.device = torch.device(env.device) if not isinstance(env.device, torch.device) else env.device
        self.model = model.to(self.device)
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        if self.world > 1:
            for p in list(self.model.parameters()) + list(self.model.buffers()):
                dist.broadcast(p.data, src=0)          # policy weights: rank 0 -> all
        self.target = copy.deepcopy(self.model).eval()
        for p in self.target.parameters():
            p.requires_grad_(False)
        on_cuda = self.device.type == "cuda"
        self.opt = torch.optim.Adam(self.model.parameters(), lr=self.cfg.lr, fused=True) if on_cuda else \
            torch.optim.Adam(self.model.parameters(), lr=self.cfg.lr)
        # learner step as a CUDA graph (the Q-network is ~600 small launches per update: launch bound at batch 512 / world)
        self.use_graph = on_cuda and self.cfg.cuda_graph
        self._graph = None
        self._static = None
        self._graph_warmup = 0
        self.replay = DeviceReplay(max(self.cfg.replay_capacity // self.world, env.B), env.n_cells, self.device)
        self.gen = torch.Generator(device=self.device)
        self.gen.manual_seed(self.cfg.seed + 1000 * self.rank)
        self.timesteps = 0            # global env steps (all ranks)
        self.updates = 0
        self.last_target_sync = 0
        self.learn_credit_from = None   # timestep at which learning started (updates owed are counted from here)
        self._updates_at_credit = 0     # learner updates already done when the credit started (non-zero after a resume)
        self._resumed = False           # a full checkpoint was loaded: train() must not re-seed the RNGs it restored
        self.log_dir = log_dir
        self.obs = None
        self.max_efficiency = float("-inf")
        # best structure over finished AND running episodes of this rank (the env resets its per-episode best on auto-reset;
        # the reference dumps the best structure at every episode end, main.py:76-85)
        self.best_eff = float("-inf")
        self.best_pattern = None
        self.backlog_warned = False
        self.timers = {"env": 0.0, "act": 0.0, "learn": 0.0}

    # --------------------------------------------------------------------------------------------
    @torch.no_grad()
    def act(self, obs, epsilon: float):
        self.model.eval()
        q = self.model(obs.to(torch.float32))
        greedy = q.argmax(dim=1)
        rnd = torch.randint(0, q.shape[1], greedy.shape, device=q.device, generator=self.gen)
        explore = torch.rand(greedy.shape, device=q.device, generator=self.gen) < epsilon
        return torch.where(explore, rnd, greedy)

    def _loss(self, obs, act, rew, nxt, done):
        cfg = self.cfg
        q = self.model(obs).gather(1, act[:, None]).squeeze(1)
        with torch.no_grad():
            qn = self.target(nxt).max(dim=1).values
            tgt = rew + cfg.gamma * (1.0 - done) * qn
        return nn.functional.huber_loss(q, tgt)

    def _allreduce_grads(self):
        if self.world == 1:
            return
        grads = [p.grad for p in self.model.parameters() if p.grad is not None]
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat)
        flat /= self.world
        torch._foreach_copy_(grads, [f.view_as(g) for f, g in zip(flat.split([g.numel() for g in grads]), grads)])

    def learn(self):
        cfg = self.cfg
        b = max(cfg.train_batch_size // self.world, 1)
        obs, act, rew, nxt, done = self.replay.sample(b, self.gen)
        obs, nxt, done = obs.to(torch.float32), nxt.to(torch.float32), done.to(torch.float32)
        self.model.train()
        if not self.use_graph:
            loss = self._loss(obs, act, rew, nxt, done)
            self.opt.zero_grad(set_to_none=True)
            loss.backward()
            self._allreduce_grads()
            self.opt.step()
            self.updates += 1
            return loss.detach()
        # ---- CUDA-graph path: static input tensors, forward + backward captured once (after three eager warm-up updates on a
        #      side stream, as torch.cuda.graphs requires); gradient all_reduce and the fused Adam step stay outside the graph ----
        if self._static is None:
            self._static = [torch.empty_like(t) for t in (obs, act, rew, nxt, done)]
        for st_, src in zip(self._static, (obs, act, rew, nxt, done)):
            st_.copy_(src)
        if self._graph is None:
            if self._graph_warmup < 3:
                side = torch.cuda.Stream(self.device)
                side.wait_stream(torch.cuda.current_stream(self.device))
                with torch.cuda.stream(side):
                    self.opt.zero_grad(set_to_none=True)
                    loss = self._loss(*self._static)
                    loss.backward()
                torch.cuda.current_stream(self.device).wait_stream(side)
                self._graph_warmup += 1
                self._allreduce_grads()
                self.opt.step()
                self.updates += 1
                return loss.detach()
            self.opt.zero_grad(set_to_none=True)
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._static_loss = self._loss(*self._static)
                self._static_loss.backward()
        self._graph.replay()
        self._allreduce_grads()
        self.opt.step()
        self.updates += 1
        return self._static_loss.detach().clone()

    @torch.no_grad()
    def sync_target(self):
        tau = self.cfg.tau
        for pt, p in zip(self.target.parameters(), self.model.parameters()):
            pt.mul_(1.0 - tau).add_(p, alpha=tau)
        for bt, b in zip(self.target.buffers(), self.model.buffers()):
            bt.copy_(b)

    # --------------------------------------------------------------------------------------------
    def train(self, total_timesteps: int, log_every: int = 0):
        """Runs until ``total_timesteps`` GLOBAL env steps have been taken.  Returns a metrics dict."""
        env, cfg = self.env, self.cfg
        if self.obs is None:
            if not self._resumed:         # fresh start only: a resumed run continues the RNG streams of its checkpoint
                np.random.seed(cfg.seed + self.rank)
                torch.manual_seed(cfg.seed + self.rank)
            self.obs = env.reset()
        per_iter = env.B * self.world
        losses = []
        sync = (lambda: torch.cuda.synchronize(self.device)) if self.device.type == "cuda" else (lambda: None)
        t_start = time.perf_counter()
        it = 0
        while self.timesteps < total_timesteps:
            eps = epsilon_at(cfg, self.timesteps)
            t0 = time.perf_counter()
            actions = self.act(self.obs, eps)
            sync(); t1 = time.perf_counter()
            nxt, reward, done, info = env.step(actions)
            # with auto-reset the returned obs of a finished env is the first obs of the next episode; the transition stored
            # is (s, a, r, s', done) with s' = the flipped pattern, reconstructed from s and a
            s_next = self.obs.clone().to(torch.int8)
            inb = (actions >= 0) & (actions < env.n_cells)
            rows = torch.nonzero(inb).squeeze(1)
            s_next[rows, actions[rows]] = -s_next[rows, actions[rows]]
            self.replay.add(self.obs, actions, reward, s_next, done.to(torch.bool))
            self.obs = nxt
            if isinstance(info, dict) and "episode_best_eff" in info and info["episode_best_eff"].numel():
                self._fold_best(info["episode_best_eff"], info["episode_best_patterns"])
            sync(); t2 = time.perf_counter()
            self.timesteps += per_iter
            if self.replay.size * self.world >= cfg.learning_starts:
                if self.learn_credit_from is None:
                    self.learn_credit_from = self.timesteps - per_iter
                    self._updates_at_credit = self.updates
                # one learner update per `timesteps_per_update` sampled timesteps (configs/simple_q.py:71-73), never more
                owed = ((self.timesteps - self.learn_credit_from) // max(cfg.timesteps_per_update, 1)
                        - (self.updates - self._updates_at_credit))
                if owed > cfg.max_updates_per_iter and not self.backlog_warned and self.rank == 0:
                    self.backlog_warned = True
                    print(f"[simple_q] {per_iter} timesteps per iteration owe {owed} learner updates, more than "
                          f"max_updates_per_iter={cfg.max_updates_per_iter}: the sample-to-update ratio of the reference "
                          f"is not reached (raise max_updates_per_iter or train_batch_size)", flush=True)
                for _ in range(int(min(max(owed, 0), cfg.max_updates_per_iter))):
                    losses.append(self.learn())
                if self.timesteps - self.last_target_sync >= cfg.target_network_update_freq:
                    self.sync_target()
                    self.last_target_sync = self.timesteps
            sync(); t3 = time.perf_counter()
            self.timers["act"] += t1 - t0; self.timers["env"] += t2 - t1; self.timers["learn"] += t3 - t2
            it += 1
            if log_every and it % log_every == 0 and self.rank == 0:
                print(f"[simple_q] t={self.timesteps} eps={eps:.3f} max_eff={max(self.best_eff, float(env.best_eff.max())):.4f} "
                      f"loss={float(losses[-1]) if losses else float('nan'):.3e}", flush=True)
        wall = time.perf_counter() - t_start
        self._fold_best(env.best_eff, env.best_patterns)          # episodes still running
        best = torch.tensor([self.best_eff], dtype=torch.float64, device=self.device)
        if self.world > 1:
            dist.all_reduce(best, op=dist.ReduceOp.MAX)
        self.max_efficiency = float(best)
        total = sum(self.timers.values()) or 1.0
        return {"timesteps": self.timesteps, "updates": self.updates, "max_efficiency": self.max_efficiency,
                "env_steps_per_s": it * per_iter / wall, "wall_s": wall,
                "time_fraction": {k: v / total for k, v in self.timers.items()},
                "final_loss": float(losses[-1]) if losses else None, "epsilon": epsilon_at(cfg, self.timesteps)}

    def _fold_best(self, effs, patterns):
        i = int(effs.argmax())
        e = float(effs[i])
        if e > self.best_eff:
            self.best_eff = e
            self.best_pattern = patterns[i].to(torch.float64).cpu().numpy().copy()

    def save_best(self):
        """main.py:80-85: dump the best structure seen by this rank (finished and running episodes) as .npy, file name from
        eff * 100."""
        if not self.log_dir:
            return None
        os.makedirs(self.log_dir, exist_ok=True)
        self._fold_best(self.env.best_eff, self.env.best_patterns)
        name = "w" + str(self.rank) + f"_{self.best_eff * 100:.6f}".replace(".", "-")
        path = os.path.join(self.log_dir, name)
        np.save(path, self.best_pattern)
        return path + ".npy"

    # --------------------------------------------------------------------------------------------
    def save_checkpoint(self, path: str):
        """Policy, target, optimiser, counters and RNG state (the reference checkpoints through Ray AIR every iteration and
        at the end, main.py:176-182).  Every rank holds the same policy; rank 0 writes.  COLLECTIVE when world > 1: every rank
        must call it, because the exploration / replay generators differ per rank and are gathered into rank 0's file."""
        rng = {"gen_state": self.gen.get_state().cpu(), "torch_rng": torch.get_rng_state(), "numpy_rng": np.random.get_state()}
        rng_per_rank = [rng]
        if self.world > 1:
            rng_per_rank = [None] * self.world
            dist.all_gather_object(rng_per_rank, rng)
        if self.rank != 0:
            return None
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        state = {"model": self.model.state_dict(), "target": self.target.state_dict(), "optimizer": self.opt.state_dict(),
                 "timesteps": self.timesteps, "updates": self.updates, "last_target_sync": self.last_target_sync,
                 "learn_credit_from": self.learn_credit_from, "best_eff": self.best_eff, "best_pattern": self.best_pattern,
                 "gen_state": rng["gen_state"], "torch_rng": rng["torch_rng"], "numpy_rng": rng["numpy_rng"],
                 "rng_per_rank": rng_per_rank, "config": dataclasses.asdict(self.cfg)}
        tmp = path + ".tmp"
        torch.save(state, tmp)
        os.replace(tmp, path)
        return path

    def load_checkpoint(self, path: str, weights_only: bool = False):
        """Resume (``--transfer_ckpt`` of the reference, main.py:163-164).  ``weights_only`` loads the two networks and leaves
        optimiser, counters and exploration schedule fresh (transfer to another wavelength / angle)."""
        state = torch.load(path, map_location=self.device, weights_only=False)
        self.model.load_state_dict(state["model"])
        self.target.load_state_dict(state.get("target", state["model"]))
        if weights_only:
            return state
        self.opt.load_state_dict(state["optimizer"])
        self.timesteps, self.updates = int(state["timesteps"]), int(state["updates"])
        self.last_target_sync = state["last_target_sync"]
        # the replay buffer is not in the checkpoint: the update credit restarts when the refilled buffer reaches
        # learning_starts again (otherwise the timesteps spent refilling would be paid back as one burst of updates on a
        # nearly empty buffer, above the reference's one update per timesteps_per_update)
        self.learn_credit_from = None
        self.best_eff, self.best_pattern = state["best_eff"], state["best_pattern"]
        per_rank = state.get("rng_per_rank") or [state]
        if len(per_rank) == self.world:                   # same world size: every rank continues its own streams
            mine = per_rank[self.rank]
            self.gen.set_state(mine["gen_state"].cpu())
            torch.set_rng_state(mine["torch_rng"].cpu())
            np.random.set_state(mine["numpy_rng"])
        else:                                             # world size changed: fresh, distinct, checkpoint-dependent streams
            salt = self.cfg.seed + 1000 * self.rank + self.timesteps
            self.gen.manual_seed(salt)
            torch.manual_seed(salt + 1)
            np.random.seed((salt + 2) % (2 ** 32))
        self._resumed = True
        self._graph = None                # the captured learner graph holds the old parameter values' addresses: still valid,
        self._graph_warmup = 0            # but re-capture keeps the optimiser state and the graph consistent
        return state
