"""GPU (B200): the native SimpleQ loop on the real vectorised env -- CUDA-graph learner path on,
64 envs: finite loss, weights move, graph replay equals the eager step on one batch, best structure survives the auto-reset,
checkpoint round trip.  RL-loop parity with RLlib is unpinned (its internals are not in the reference tree); these are
smoke / consistency checks of the device path that the CPU tests cannot reach."""
import copy

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_simple_q_on_vector_env_cuda_graph(tmp_path):
    import torch
    from physics_informed_reinforcement_learning_b200 import VectorDeflectorEnv
    from physics_informed_reinforcement_learning_b200.qnet import ShallowUQNet
    from physics_informed_reinforcement_learning_b200.simple_q import SimpleQConfig, SimpleQTrainer
    torch.cuda.set_device(0)
    torch.manual_seed(0)
    env = VectorDeflectorEnv(64, max_episode_steps=16, device=0, data_dir=str(tmp_path), check_status_every=8)
    cfg = SimpleQConfig(learning_starts=256, train_batch_size=128, epsilon_timesteps=2000, target_network_update_freq=512,
                        replay_capacity=8192, cuda_graph=True)
    model = ShallowUQNet(256)
    w0 = copy.deepcopy(model.state_dict())
    tr = SimpleQTrainer(env, model, cfg, log_dir=str(tmp_path))
    out = tr.train(2048)
    assert out["timesteps"] >= 2048 and out["updates"] >= 40
    assert tr._graph is not None                                     # the captured learner step was used
    assert np.isfinite(out["final_loss"])
    moved = max(float((model.state_dict()[k].float() - w0[k].float().cuda()).abs().max()) for k in w0 if "weight" in k)
    assert moved > 0.0
    # graph replay == eager forward + backward on the same static batch (the graph accumulates into the .grad buffers it captured)
    params = list(tr.model.parameters())
    tr.model.train()
    tr.opt.zero_grad(set_to_none=False)
    tr._graph.replay()
    torch.cuda.synchronize()
    g_graph = [p.grad.clone() for p in params]
    loss_g = float(tr._static_loss.detach())
    loss_e = tr._loss(*tr._static)
    g_eager = torch.autograd.grad(loss_e, params)
    assert abs(loss_g - float(loss_e.detach())) <= 1e-5 * max(1.0, abs(loss_g))
    for ge, gg in zip(g_eager, g_graph):
        assert torch.allclose(ge, gg, rtol=1e-3, atol=1e-6)
    # 2048 steps / 64 envs = 32 iterations = two 16-step episodes per env: every env was auto-reset once
    assert env.episodes_done >= 64
    eff, pat = env.best()
    assert eff >= float(env.best_eff.max()) - 1e-15 and abs(out["max_efficiency"] - eff) <= 1e-12
    assert abs(float(env.get_efficiency_batch(pat[None, :])[0]) - eff) <= 1e-12      # the stored structure reproduces its efficiency
    import os
    assert len(os.listdir(os.path.join(str(tmp_path), "MeentIndex"))) >= 128          # one dump per reset (64 initial + 64 auto)
    ck = tr.save_checkpoint(str(tmp_path / "ck.pt"))
    tr2 = SimpleQTrainer(VectorDeflectorEnv(8, max_episode_steps=16, device=0), ShallowUQNet(256), cfg)
    tr2.load_checkpoint(ck)
    assert tr2.timesteps == tr.timesteps and all(torch.equal(a, b) for a, b in zip(tr.model.parameters(), tr2.model.parameters()))
