"""CPU: the Q-network hook and the native SimpleQ loop -- architecture compatibility with the
reference's pretrained state_dict layout, the reference's hyper-parameter schedule, replay ring semantics, and a two-rank
gloo run that must keep the policy weights identical across ranks."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from physics_informed_reinforcement_learning_b200.qnet import ShallowUQNet, FCNQNet
from physics_informed_reinforcement_learning_b200.simple_q import SimpleQConfig, SimpleQTrainer, DeviceReplay, epsilon_at


class ToyEnv:
    """VectorDeflectorEnv's interface with a closed-form 'efficiency' (fraction of cells equal to a target), CPU only."""

    def __init__(self, B, n=16, horizon=8, seed=0):
        self.B, self.n_cells, self.device = B, n, torch.device("cpu")
        self.h = horizon
        g = torch.Generator().manual_seed(seed)
        self.target = (torch.randint(0, 2, (n,), generator=g) * 2 - 1).to(torch.int8)
        self.patterns = torch.ones(B, n, dtype=torch.int8)
        self.best_patterns = self.patterns.clone()
        self.elapsed = torch.zeros(B, dtype=torch.int64)

    def _eff(self):
        return (self.patterns == self.target).to(torch.float64).mean(dim=1)

    def reset(self):
        self.patterns.fill_(1)
        self.eff = self._eff()
        self.best_eff = self.eff.clone()
        self.elapsed.zero_()
        return self.patterns.to(torch.float64)

    def step(self, actions):
        rows = torch.arange(self.B)
        self.patterns[rows, actions] = -self.patterns[rows, actions]
        new = self._eff()
        reward = new - self.eff
        self.eff = new
        better = new > self.best_eff
        self.best_eff = torch.where(better, new, self.best_eff)
        self.best_patterns[better] = self.patterns[better]
        self.elapsed += 1
        done = self.elapsed >= self.h
        info = {}
        if done.any():
            # same contract as VectorDeflectorEnv.step: episode results go to `info`, then the per-episode best is re-seeded
            info = {"episode_best_eff": self.best_eff[done].clone(), "episode_best_patterns": self.best_patterns[done].clone()}
            self.patterns[done] = 1
            self.eff = self._eff()
            self.best_eff[done] = self.eff[done]
            self.best_patterns[done] = self.patterns[done]
            self.elapsed[done] = 0
        return self.patterns.to(torch.float64), reward, done, info


def test_unet_matches_reference_layout():
    m = ShallowUQNet(256)
    assert sum(p.numel() for p in m.parameters()) == 1_974_113
    keys = set(m.state_dict().keys())
    for k in ("conv1_1.weight", "conv1_2.convrelu.0.weight", "conv1_2.convrelu.1.running_mean", "conv6_6.convrelu.1.bias",
              "conv8_1.weight", "conv11_6.convrelu.0.bias", "conv11_fin.weight"):
        assert k in keys, k
    assert m.state_dict()["conv8_1.weight"].shape == (128, 384, 3)
    assert m.state_dict()["conv11_1.weight"].shape == (16, 48, 3)
    m.eval()
    x = torch.sign(torch.randn(3, 1, 256))
    q = m(x)
    assert q.shape == (3, 256)
    assert torch.allclose(m(x[:, 0]), q)                                  # (B, n) and (B, 1, n) observations
    # circular convolutions + pooling by 16: translating the pattern by 16 cells translates the Q-values
    assert torch.allclose(m(torch.roll(x, 16, dims=2)), torch.roll(q, 16, dims=1), atol=1e-5)
    assert FCNQNet(256)(x).shape == (3, 256) and FCNQNet(256, 1200).fc2.weight.shape == (1200, 1200)


def test_schedule_and_replay():
    cfg = SimpleQConfig()
    assert epsilon_at(cfg, 0) == pytest.approx(0.99) and epsilon_at(cfg, 100_000) == pytest.approx(0.01)
    assert epsilon_at(cfg, 50_000) == pytest.approx(0.5) and epsilon_at(cfg, 10 ** 7) == pytest.approx(0.01)
    assert (cfg.gamma, cfg.lr, cfg.train_batch_size, cfg.replay_capacity, cfg.learning_starts,
            cfg.target_network_update_freq, cfg.tau) == (0.99, 1e-3, 512, 100_000, 1000, 2000, 0.1)
    r = DeviceReplay(10, 4, "cpu")
    for k in range(4):
        o = torch.full((3, 4), k, dtype=torch.int8)
        r.add(o, torch.full((3,), k), torch.full((3,), float(k)), o + 1, torch.zeros(3, dtype=torch.bool))
    assert r.size == 10 and r.pos == 2
    assert r.obs[0, 0] == 3 and r.obs[1, 0] == 3 and r.obs[2, 0] == 0     # ring wrapped: slots 0,1 hold the newest rows
    o, a, rew, n, d = r.sample(64)
    assert bool((n == o + 1).all()) and bool((a.to(torch.float32) == rew).all())


def test_trainer_learns_toy_env():
    torch.manual_seed(0)
    env = ToyEnv(32, n=16, horizon=8)
    cfg = SimpleQConfig(learning_starts=64, train_batch_size=64, epsilon_timesteps=3000, target_network_update_freq=200,
                        replay_capacity=4096)
    tr = SimpleQTrainer(env, FCNQNet(16, 64), cfg)
    out = tr.train(6000)
    assert out["updates"] > 100 and out["timesteps"] >= 6000
    assert out["max_efficiency"] >= 0.9        # the toy optimum (1.0) is reachable within 8 flips from all-ones
    # greedy policy after training beats a random policy
    obs = env.reset()
    tot = torch.zeros(32, dtype=torch.float64)
    for _ in range(8):
        obs, r, d, _ = env.step(tr.act(obs, 0.0))
        tot += r
    assert float(tot.mean()) > 0.1


def test_best_survives_auto_reset_and_checkpoint_roundtrip(tmp_path):
    """max_efficiency must not fall when episodes end (the env re-seeds its per-episode best on auto-reset), and a
    checkpoint restores policy, optimiser, counters and the best structure."""
    torch.manual_seed(1)
    env = ToyEnv(8, n=16, horizon=8)
    cfg = SimpleQConfig(learning_starts=32, train_batch_size=32, epsilon_timesteps=400, target_network_update_freq=64,
                        replay_capacity=1024)
    tr = SimpleQTrainer(env, FCNQNet(16, 32), cfg, log_dir=str(tmp_path))
    seen = []
    for stop in (8 * 8 * 3, 8 * 8 * 6, 8 * 8 * 6 + 8):          # ... the last stop ends right after a reset of every env
        seen.append(tr.train(stop)["max_efficiency"])
    assert seen[0] <= seen[1] <= seen[2] and seen[2] > float(env.best_eff.max()) - 1e-12
    assert tr.best_pattern is not None and tr.best_pattern.shape == (16,)
    best_file = tr.save_best()
    assert np.array_equal(np.load(best_file), tr.best_pattern)
    ck = tr.save_checkpoint(str(tmp_path / "ckpt.pt"))
    tr2 = SimpleQTrainer(ToyEnv(8, n=16, horizon=8), FCNQNet(16, 32), cfg)
    tr2.load_checkpoint(ck)
    assert (tr2.timesteps, tr2.updates, tr2.best_eff) == (tr.timesteps, tr.updates, tr.best_eff)
    for a, b in zip(tr.model.parameters(), tr2.model.parameters()):
        assert torch.equal(a, b)
    assert tr2.opt.state_dict()["state"][0]["exp_avg"].equal(tr.opt.state_dict()["state"][0]["exp_avg"])
    tr3 = SimpleQTrainer(ToyEnv(8, n=16, horizon=8), FCNQNet(16, 32), cfg)
    tr3.load_checkpoint(ck, weights_only=True)
    assert tr3.timesteps == 0 and all(torch.equal(a, b) for a, b in zip(tr.model.parameters(), tr3.model.parameters()))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(100 + rank)                     # different initial weights: the broadcast must align them
    env = ToyEnv(8, n=16, horizon=8, seed=0)
    cfg = SimpleQConfig(learning_starts=32, train_batch_size=32, epsilon_timesteps=500, target_network_update_freq=64,
                        replay_capacity=1024)
    tr = SimpleQTrainer(env, FCNQNet(16, 32), cfg)
    out = tr.train(800)
    flat = torch.cat([p.detach().reshape(-1) for p in tr.model.parameters()])
    gathered = [torch.zeros_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    q.put((rank, out["timesteps"], out["updates"], float((gathered[0] - gathered[1]).abs().max()), out["max_efficiency"]))
    dist.destroy_process_group()


def test_two_rank_gloo_keeps_weights_in_sync():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29650 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(30)
    assert all(p.exitcode == 0 for p in procs)
    for rank, ts, upd, diff, mx in res:
        assert ts >= 800 and upd > 10
        assert diff <= 1e-6                            # gradient all_reduce keeps every rank's policy identical
    assert res[0][4] == res[1][4]                      # max_efficiency reduced with MAX over ranks


def _explore_draw(tr):
    """Next exploration actions of a trainer at epsilon = 1: they depend on its private generator only, not on the policy."""
    return tr.act(torch.ones(tr.env.B, tr.env.n_cells, dtype=torch.float64), 1.0).clone()


def test_resume_continues_rng_streams_and_rebases_update_credit(tmp_path):
    """A resumed run must continue the RNG streams stored in the checkpoint (train() used to re-seed them), and must not pay
    back the timesteps spent refilling the empty replay buffer as one burst of learner updates."""
    cfg = SimpleQConfig(learning_starts=64, train_batch_size=16, epsilon_timesteps=400, target_network_update_freq=64,
                        replay_capacity=1024, timesteps_per_update=8, max_updates_per_iter=64)
    tr = SimpleQTrainer(ToyEnv(8, n=16, horizon=8), FCNQNet(16, 32), cfg)
    tr.train(8 * 20)
    ck = tr.save_checkpoint(str(tmp_path / "ckpt.pt"))
    np_next, torch_next = np.random.random(), torch.rand(1)          # what the uninterrupted run would draw next
    want = _explore_draw(tr)
    np.random.seed(999); torch.manual_seed(999)                       # the resuming process starts from unrelated RNG state
    tr2 = SimpleQTrainer(ToyEnv(8, n=16, horizon=8), FCNQNet(16, 32), cfg)
    tr2.load_checkpoint(ck)
    tr2.train(tr2.timesteps)                                          # runs train()'s start-up code (reset), no iteration
    assert np.random.random() == np_next and torch.equal(torch.rand(1), torch_next)
    assert torch.equal(_explore_draw(tr2), want)
    # update credit: while the buffer refills (64 transitions = 8 iterations) nothing is owed; afterwards one update per
    # timesteps_per_update = one per iteration of 8 envs, never a catch-up burst
    u0 = tr2.updates
    tr2.train(tr2.timesteps + 8 * 8)
    assert tr2.updates - u0 <= 1
    tr2.train(tr2.timesteps + 8 * 5)
    assert 4 <= tr2.updates - u0 <= 6


def _resume_worker(rank, world, port, q, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = SimpleQConfig(learning_starts=32, train_batch_size=32, epsilon_timesteps=500, target_network_update_freq=64,
                        replay_capacity=1024)
    tr = SimpleQTrainer(ToyEnv(8, n=16, horizon=8, seed=0), FCNQNet(16, 32), cfg)
    tr.train(160)
    path = os.path.join(tmp, "ckpt.pt")
    tr.save_checkpoint(path)                          # collective: every rank's generator goes into rank 0's file
    dist.barrier()
    want = _explore_draw(tr)
    tr2 = SimpleQTrainer(ToyEnv(8, n=16, horizon=8, seed=0), FCNQNet(16, 32), cfg)
    tr2.load_checkpoint(path)
    tr2.train(tr2.timesteps)
    got = _explore_draw(tr2)
    q.put((rank, got.tolist(), want.tolist()))
    dist.destroy_process_group()


def test_two_rank_resume_keeps_per_rank_exploration_streams(tmp_path):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29850 + os.getpid() % 100
    procs = [ctx.Process(target=_resume_worker, args=(r, 2, port, q, str(tmp_path))) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(30)
    assert all(p.exitcode == 0 for p in procs)
    for rank, got, want in res:
        assert got == want                             # each rank continues ITS stream, as an uninterrupted run would
    assert res[0][1] != res[1][1]                      # and the ranks do not share one stream after the resume
