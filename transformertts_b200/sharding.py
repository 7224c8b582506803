"""Batch sharding for multi-GPU runs: utterances are independent, so a batch splits into contiguous
blocks of utterances, one block per rank, with NO data-path collective.  Rows arrive sorted by phoneme length
(dataset.py:74-76), so contiguous blocks are also length-bucketed.  Only the optional gather of results and the
timing barrier use torch.distributed."""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch


def shard_bounds(n_items: int, world_size: int, rank: int) -> Tuple[int, int]:
    """[lo, hi) of the contiguous block owned by `rank`; sizes differ by at most one, earlier ranks get the extra."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad world_size / rank")
    base, extra = divmod(n_items, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def micro_batches(lo: int, hi: int, size: int) -> List[Tuple[int, int]]:
    """Split [lo, hi) into consecutive micro-batches of at most `size` utterances (each one `inference` call:
    the reference's all-rows stop rule, model/model.py:382-384, couples the rows of ONE call only)."""
    if size <= 0:
        raise ValueError("micro-batch size must be positive")
    return [(s, min(s + size, hi)) for s in range(lo, hi, size)]


def sharded_inference(model, phoneme: torch.Tensor, phoneme_lens: torch.Tensor, micro_batch: int = 32,
                      max_len: int = 1500, stop_threshold: float = 0.5, rank: Optional[int] = None,
                      world_size: Optional[int] = None, gather: bool = False):
    """Run `model.inference` on this rank's block of utterances, micro-batch by micro-batch.
    Returns a list of (lo, hi, result_dict); with gather=True every rank receives all ranks' lists
    (torch.distributed.all_gather_object -- control plane only, after the decode has finished)."""
    import torch.distributed as dist
    if world_size is None:
        world_size = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    lo, hi = shard_bounds(phoneme.shape[0], world_size, rank)
    out = []
    for s, e in micro_batches(lo, hi, micro_batch):
        tp = int(phoneme_lens[s:e].max())                     # the reference pads a batch to its own longest row
        res = model.inference(phoneme[s:e, :tp], phoneme_lens[s:e], max_len=max_len, stop_threshold=stop_threshold)
        out.append((s, e, res))
    if gather and world_size > 1:
        host = [(s, e, {k: v.cpu() for k, v in r.items()}) for s, e, r in out]
        all_parts: List = [None] * world_size
        dist.all_gather_object(all_parts, host)
        out = [x for part in all_parts for x in part]
    return out
