"""Vocabulary-sharded generation (BASELINE.json config 4): the decoder rows are split
over ranks; every rank keeps the (cheap) recurrence replicated, decodes its own slice with the fused
decoder+race kernel, and ONE all-gather of (key, id, banned keys) per token lets all ranks pick the
same token.  Because the sampling race is a max over independent keys, the merge is exact: the ids
equal what a single unsharded engine draws with the same seed.

`torch.distributed` (NCCL) is the only collective used anywhere in this package, and only here.
With `logical_shards=n` the same code runs n slices on one GPU with no process group (tests).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Tuple

import torch

from . import _lib
from .engine import Engine, SamplingParams


def shard_bounds(vocab_sz: int, n: int) -> List[Tuple[int, int]]:
    """n contiguous slices made of whole 256-row decoder tiles, as even as possible."""
    tiles = -(-vocab_sz // 256)
    if tiles < n:
        raise ValueError(f'a {vocab_sz}-row vocabulary has only {tiles} decoder tiles for {n} shards')
    out, t0 = [], 0
    for r in range(n):
        t1 = t0 + tiles // n + (1 if r < tiles % n else 0)
        out.append((t0 * 256, min(t1 * 256, vocab_sz)))
        t0 = t1
    return out


def generate_vocab_sharded(engine: Engine, prompt: torch.Tensor, n_new: int, sampling: SamplingParams,
                           group=None, logical_shards: Optional[int] = None) -> torch.Tensor:
    """prompt int64 (P,bs), identical on every rank -> ids (n_new,bs), identical on every rank."""
    import torch.distributed as dist
    lib, dev = engine.lib, engine.device
    if logical_shards is None:
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        my = [shard_bounds(engine.vocab_sz, world)[rank]]
    else:
        world, rank = logical_shards, 0
        my = shard_bounds(engine.vocab_sz, world)
    P, bs = prompt.shape
    st = engine._stream()
    s, keep = sampling.to_c()
    engine.reset(bs)
    engine.forward(prompt, need_logits=False, need_raw=False)          # prompt, replicated
    keys = torch.empty(world, bs, dtype=torch.float32, device=dev)
    ids = torch.empty(world, bs, dtype=torch.int64, device=dev)
    banned = torch.empty(world, bs, 64, dtype=torch.float32, device=dev)
    out = torch.empty(n_new, bs, dtype=torch.int64, device=dev)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    for i in range(n_new):
        for j, (v0, v1) in enumerate(my):
            slot = j if logical_shards is not None else rank
            _lib.check(lib.awd_decode_race_shard(engine._h, v0, v1, C.byref(s), i, ptr(keys[slot]), ptr(ids[slot]),
                                                 ptr(banned[slot]), st))
        if logical_shards is None and world > 1:       # 8 + 4*64 bytes per stream and rank
            dist.all_gather_into_tensor(keys, keys[rank].clone(), group=group)
            dist.all_gather_into_tensor(ids, ids[rank].clone(), group=group)
            dist.all_gather_into_tensor(banned, banned[rank].clone(), group=group)
        _lib.check(lib.awd_merge_race(ptr(keys), ptr(ids), ptr(banned), world, bs, engine.vocab_sz, C.byref(s), ptr(out[i]), st))
        engine.forward(out[i:i + 1], need_logits=False, need_raw=False)   # feed back (main.py:71), replicated
    return out
