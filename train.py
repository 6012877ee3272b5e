#!/usr/bin/env python
"""RL training entry point: the reference's ``python main.py`` (main.py:88-187) over the B200-native vectorised env.

    python train.py --wavelength 1100 --angle 60 --thickness 325 --train_steps 200000 --num_envs 1024
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 train.py --num_envs 1024

``--num_envs`` is the GLOBAL number of vectorised envs (config C3: 1024 = 128 per GPU x 8); each rank owns num_envs / world
of them, its own RCWA handle and replay shard.  The Q-network is the unchanged UNet hook (qnet.ShallowUQNet); a pretrained
state_dict is loaded when ``--pretrained_ckpt`` exists (the reference's file is not shipped: .MISSING_LARGE_BLOBS).
"""
from __future__ import annotations

import argparse
import json
import os

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--wavelength", type=int, default=1100)
    ap.add_argument("--angle", type=int, default=60)
    ap.add_argument("--thickness", type=int, default=325)
    ap.add_argument("--order", type=int, default=40)
    ap.add_argument("--n_cells", type=int, default=256)
    ap.add_argument("--train_steps", type=int, default=200000)
    ap.add_argument("--num_envs", type=int, default=1024)
    ap.add_argument("--max_episode_steps", type=int, default=128)
    ap.add_argument("--pretrained_ckpt", type=str, default="pretrained/pretrained_1100nm_60degree.pt")
    ap.add_argument("--model", choices=["unet", "fcn", "fcn_heavy"], default="unet")
    ap.add_argument("--data_dir", type=str, default="runs/simple_q")
    ap.add_argument("--log_every", type=int, default=50)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--no_cuda_graph", action="store_true", help="run the learner step eagerly")
    ap.add_argument("--transfer_ckpt", type=str, default=None,
                    help="checkpoint written by this script to resume from (main.py:163-164 algo.load_checkpoint): networks, optimiser, "
                         "counters and every rank's RNG streams; the replay buffer is NOT stored, so learning pauses until "
                         "the refilled buffer reaches learning_starts and the update credit restarts there")
    ap.add_argument("--transfer_weights_only", action="store_true",
                    help="with --transfer_ckpt: load the networks only (fresh optimiser, counters and exploration schedule)")
    ap.add_argument("--checkpoint_every", type=int, default=20000,
                    help="global timesteps between checkpoints <data_dir>/checkpoint.pt (0: only at the end)")
    ap.add_argument("--write_structures", action="store_true",
                    help="StructureWriter (utils.py:36-82): dump every reset structure as .npy under <data_dir>/MeentIndex")
    ap.add_argument("--timesteps_per_update", type=int, default=32,
                    help="sampled env steps per learner update (reference: 16 workers x fragment length 2)")
    args = ap.parse_args()

    from physics_informed_reinforcement_learning_b200 import VectorDeflectorEnv
    from physics_informed_reinforcement_learning_b200.qnet import ShallowUQNet, FCNQNet, load_pretrained
    from physics_informed_reinforcement_learning_b200.simple_q import SimpleQConfig, SimpleQTrainer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank = dist.get_rank() if world > 1 else 0
    per_rank = max(args.num_envs // world, 1)
    env = VectorDeflectorEnv(per_rank, n_cells=args.n_cells, wavelength=args.wavelength, desired_angle=args.angle,
                             order=args.order, thickness=args.thickness, max_episode_steps=args.max_episode_steps, device=local,
                             data_dir=args.data_dir if args.write_structures else None, check_status_every=64)
    model = {"unet": lambda: ShallowUQNet(args.n_cells), "fcn": lambda: FCNQNet(args.n_cells),
             "fcn_heavy": lambda: FCNQNet(args.n_cells, 1200)}[args.model]()
    loaded = load_pretrained(model, args.pretrained_ckpt)
    trainer = SimpleQTrainer(env, model, SimpleQConfig(seed=args.seed, timesteps_per_update=args.timesteps_per_update, cuda_graph=not args.no_cuda_graph), log_dir=args.data_dir)
    if args.transfer_ckpt:
        trainer.load_checkpoint(args.transfer_ckpt, weights_only=args.transfer_weights_only)
    ckpt_path = os.path.join(args.data_dir, "checkpoint.pt")
    out = {}
    while trainer.timesteps < args.train_steps:       # periodic checkpoints (Ray AIR CheckpointConfig, main.py:176-182)
        stop = args.train_steps if args.checkpoint_every <= 0 else min(args.train_steps, trainer.timesteps + args.checkpoint_every)
        out = trainer.train(stop, log_every=args.log_every)
        trainer.save_checkpoint(ckpt_path)
    env.check_status()
    path = trainer.save_best()
    if rank == 0:
        out.update({"world": world, "envs_per_rank": per_rank, "pretrained_loaded": loaded, "best_structure": path,
                    "checkpoint": ckpt_path})
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
