"""Multi-GPU sharding of the upsampling path.  Host logic only: no kernels, works with any
torch.distributed backend (NCCL on the GPUs, gloo in the CPU tests).

Clouds and patches are independent units (BN in eval mode, ActNorm with stored parameters), so

  * a batch of clouds is split into contiguous blocks, one per rank, and nothing is exchanged
    (`block_range`, `shard_files`: what bench.py and modules/discrete/upsample.py do);
  * ONE large cloud (config 3: a 5,000-pt cloud = 78 patches) is split by patches: every rank derives the same
    patch list (seed FPS and patch kNN are deterministic and cheap), runs the network on its block of patches, and
    a single all_gather of the [n_patch, 1280, 3] results (1.2 MB for 78 patches) precedes the merge FPS, which
    is sequential per cloud and therefore replicated (`patch_sharded`).

The reference has no multi-GPU path (upsample.py:27 hard-codes cuda:0); this is the "final output gather, no
collective on the hot path" of north_star.
"""
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def block_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of n units owned by `rank`; blocks of ceil(n / world), the last ones may be short or
    empty (78 patches over 8 ranks -> 10,10,10,10,10,10,10,8)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    per = -(-n // world)
    lo = min(rank * per, n)
    return lo, min(lo + per, n)


def shard_files(paths: Sequence[str], rank: int, world: int) -> List[str]:
    """Round-robin share of the sorted file list (every file goes to exactly one rank, whatever order os.walk gave)."""
    return list(sorted(paths)[rank::world]) if world > 1 else list(paths)


def world_info(group=None) -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def patch_sharded(sample_fn: Callable[[torch.Tensor], torch.Tensor], patches: torch.Tensor,
                  group: Optional[object] = None) -> torch.Tensor:
    """Apply a per-patch map `sample_fn: [b, N, C] -> [b, N2, C]` to `patches[Bp, N, C]` (identical on every rank) with
    the Bp patches split in contiguous blocks over the ranks of `group`, and return the full [Bp, N2, C] result on
    every rank (one all_gather of equal-sized, zero-padded blocks).  With no process group it is just sample_fn."""
    rank, world = world_info(group)
    if world == 1:
        return sample_fn(patches)
    Bp = patches.shape[0]
    per = -(-Bp // world)
    lo, hi = block_range(Bp, rank, world)
    # a rank with an empty block still runs the map on one patch to learn the output shape without communication
    probe = patches[lo:hi] if hi > lo else patches[:1]
    local = sample_fn(probe.contiguous())
    if hi == lo:
        local = local[:0]
    buf = local.new_zeros((per,) + tuple(local.shape[1:]))
    buf[:hi - lo] = local
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    return torch.cat(parts, 0)[:Bp]
