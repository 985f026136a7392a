"""Env-sharded data parallelism, one process per GPU (torchrun), mirroring
``cleanrl/ppo_atari_multigpu.py``: world/batch arithmetic (:166-172), seeding
(:208-211,:231) and the single flat-gradient SUM all-reduce per minibatch
(:360-374).  NCCL over NVLink on the GPU box; gloo for CPU tests of this logic."""
from __future__ import annotations

import os
import random
from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist


@dataclass
class WorldInfo:
    rank: int
    local_rank: int
    world_size: int

    @property
    def distributed(self):
        return self.world_size > 1


def world_from_env() -> WorldInfo:
    """torchrun exports RANK / LOCAL_RANK / WORLD_SIZE (ppo_atari_multigpu.py:165-166)."""
    return WorldInfo(int(os.getenv("RANK", "0")), int(os.getenv("LOCAL_RANK", "0")),
                     int(os.getenv("WORLD_SIZE", "1")))


def init_process_group(world: WorldInfo, backend: str = "nccl"):
    """ppo_atari_multigpu.py:174-175 (backend default there is gloo; NCCL here on GPUs)."""
    if world.distributed and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = torch.device("cuda", world.local_rank)
        dist.init_process_group(backend, rank=world.rank, world_size=world.world_size, **kwargs)


def shard_sizes(local_num_envs: int, num_steps: int, num_minibatches: int, world_size: int) -> dict:
    """ppo_atari_multigpu.py:167-172."""
    local_batch_size = int(local_num_envs * num_steps)
    local_minibatch_size = int(local_batch_size // num_minibatches)
    num_envs = local_num_envs * world_size
    batch_size = int(num_envs * num_steps)
    minibatch_size = int(batch_size // num_minibatches)
    return dict(local_batch_size=local_batch_size, local_minibatch_size=local_minibatch_size,
                num_envs=num_envs, batch_size=batch_size, minibatch_size=minibatch_size)


def seed_rank(seed: int, rank: int) -> int:
    """Per-rank python/numpy seed, common torch seed for identical initial weights
    (ppo_atari_multigpu.py:208-211).  Call ``reseed_sampling`` after the model is built."""
    rank_seed = seed + rank
    random.seed(rank_seed)
    np.random.seed(rank_seed)
    torch.manual_seed(rank_seed - rank)
    return rank_seed


def reseed_sampling(rank_seed: int):
    """ppo_atari_multigpu.py:231: action sampling differs per rank."""
    torch.manual_seed(rank_seed)


def allreduce_sum_(flat: torch.Tensor, group=None):
    """The one collective of the path (ppo_atari_multigpu.py:367), in place on the
    flat gradient arena -- no torch.cat / copy-back; the 1/world_size is folded into
    the clip+Adam kernel."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat
