"""Multi-GPU sharding over torch.distributed: one process per GPU, disjoint Philox counter ranges, and a
single all-reduce of the int64 tallies (NCCL over NVLink on a GPU box).

The path shards naturally -- samples are independent -- so there is no data-path collective; the only
exchange is the final tally sum.  The reference has no distributed code at all.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from .engine import Plan, shard_range


def run_sharded(plan: Plan, n_samples: int, seed: int = 0, sample_begin: int = 0, group=None,
                device: Optional[torch.device] = None, **run_kw) -> torch.Tensor:
    """Evaluate [sample_begin, sample_begin + n_samples) split over the ranks of `group`.

    Every rank returns the same int64 tensor of global tallies: the CUDA engine writes this rank's tallies straight into
    the tensor that the all-reduce then consumes (no host hop).  `device` defaults to the current CUDA device.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    begin, n = shard_range(sample_begin, n_samples, rank, world)
    nt = plan.n_tallies
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    out = torch.zeros(max(nt, 1), dtype=torch.int64, device=device)
    stream = torch.cuda.current_stream(device).cuda_stream if device.type == "cuda" else None
    index = device.index if device.index is not None else (torch.cuda.current_device() if device.type == "cuda" else 0)
    plan.run_device(out.data_ptr(), n, seed=seed, sample_begin=begin, device=index, stream=stream, **run_kw)
    if world > 1:
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out[:nt]
