"""Ciphertext-index sharding across the GPUs of one box (one process per GPU, torch.distributed plumbing).

The reference has no distributed code: its three calls loop over ciphertext indices independently
(ckks.cpp:71-83 chunks, :286-297 per-index scale/add, :187-205 per-ciphertext decrypt), and the learner sum is
*within* an index, so the path shards with no data-path collective.  Each rank owns a contiguous block of
ciphertext indices, i.e. a contiguous parameter range; the only collective is an optional all-gather of the
decoded float64 slices when the caller wants one contiguous buffer (NCCL over NVLink on GPUs, gloo in CPU tests).
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class Shard:
    rank: int
    world: int
    ct_first: int      # first ciphertext index owned by this rank (the sampler's ct_index0)
    ct_count: int      # ciphertexts owned (0 for trailing ranks of a tiny model)
    param_first: int   # first parameter index
    param_count: int   # parameters owned (the last owner gets the short tail)
    ct_per_rank: int   # common block size: ceil(n_ct / world)


def shard_range(n_params: int, batch: int, rank: int, world: int) -> Shard:
    """Contiguous block partition of ceil(n_params / batch) ciphertexts over `world` ranks."""
    if n_params <= 0 or batch <= 0 or world <= 0 or not 0 <= rank < world:
        raise ValueError("bad shard arguments")
    n_ct = -(-n_params // batch)
    per = -(-n_ct // world)
    first = min(rank * per, n_ct)
    count = min(per, n_ct - first)
    p_first = first * batch
    p_count = max(0, min(n_params, (first + count) * batch) - p_first) if count else 0
    return Shard(rank, world, first, count, p_first, p_count, per)


def gather_decoded(local, n_params: int, batch: int, group=None):
    """All-gather the ranks' decoded slices into one contiguous float64 tensor of n_params values.

    `local` is this rank's decoded slice (torch tensor, on the device the process group works on). Slices are
    padded to the common block size so one all_gather_into_tensor suffices; the padding is dropped afterwards.
    """
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sh = shard_range(n_params, batch, rank, world)
    if local.numel() != sh.param_count:
        raise ValueError(f"rank {rank}: slice has {local.numel()} values, shard owns {sh.param_count}")
    block = sh.ct_per_rank * batch
    padded = torch.zeros(block, dtype=local.dtype, device=local.device)
    padded[:sh.param_count] = local
    out = torch.empty(world * block, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    return out[:n_params]
