"""B200-native batched RCWA backend for DeflectorGym (hot path of jLabKAIST/Physics-Informed-Reinforcement-Learning).

Public surface:
    RCWABatch            batched pattern -> diffraction efficiency on one GPU (C-ABI CUDA library underneath)
    JLABCode             drop-in for the reference's solver adapter (one structure)
    MeentIndex, make     drop-in env ('MeentIndex-v0') with the reference's reset/step/get_efficiency
    VectorDeflectorEnv   B envs stepped by one batched launch
    shard_range, gather_efficiencies   multi-GPU helpers (one process per GPU, torch.distributed)
"""
from .solver import RCWABatch, JLABCode, shared_backend  # noqa: F401
from .envs import MeentIndex, VectorDeflectorEnv, make, random_bunch_init  # noqa: F401
from .materials import find_nk_index, n_silicon_real  # noqa: F401
from .dist import shard_range, gather_efficiencies, max_over_ranks  # noqa: F401

__all__ = ["RCWABatch", "JLABCode", "shared_backend", "MeentIndex", "VectorDeflectorEnv", "make",
           "random_bunch_init", "find_nk_index", "n_silicon_real", "shard_range", "gather_efficiencies"]
