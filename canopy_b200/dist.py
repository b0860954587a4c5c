"""Multi-GPU sharding over torch.distributed: one process per GPU, disjoint Philox counter ranges, and a
single all-reduce of the int64 tallies (NCCL over NVLink on a GPU box).

The path shards naturally -- samples are independent -- so there is no data-path collective; the only
exchange is the final tally sum.  The reference has no distributed code at all.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch
import torch.distributed as dist

from .engine import Plan, shard_range


def run_sharded(plan: Optional[Plan], n_samples: int, seed: int = 0, sample_begin: int = 0, group=None,
                device: Optional[torch.device] = None,
                local_run: Optional[Callable[[int, int, torch.Tensor], None]] = None, n_tallies: Optional[int] = None,
                **run_kw) -> torch.Tensor:
    """Evaluate [sample_begin, sample_begin + n_samples) split over the ranks of `group`.

    Every rank returns the same int64 tensor of global tallies.  `local_run(begin, n, out)` fills `out`
    with this rank's tallies; by default it is the CUDA engine writing straight into the tensor that the
    all-reduce then consumes (no host hop).  The hook exists so that the sharding + collective logic can be
    exercised with the gloo backend on a CPU-only box, where the engine itself refuses to run.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    begin, n = shard_range(sample_begin, n_samples, rank, world)
    nt = n_tallies if n_tallies is not None else plan.n_tallies
    if local_run is None:
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        out = torch.zeros(max(nt, 1), dtype=torch.int64, device=device)
        stream = torch.cuda.current_stream(device).cuda_stream
        plan.run_device(out.data_ptr(), n, seed=seed, sample_begin=begin, device=device.index, stream=stream, **run_kw)
    else:
        out = torch.zeros(max(nt, 1), dtype=torch.int64, device=device or "cpu")
        local_run(begin, n, out)
    if world > 1:
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out[:nt]
