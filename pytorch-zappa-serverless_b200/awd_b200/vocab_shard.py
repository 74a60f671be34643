"""Vocabulary-sharded generation (BASELINE.json config 4): the decoder rows are split
over ranks; every rank keeps the (cheap) recurrence replicated, decodes its own slice with the fused
decoder+race kernel, and ONE all-gather of (id, key, banned keys) per token -- one packed block per rank,
gathered in place -- lets all ranks pick the same token.  Because the sampling race is a max over independent keys, the merge is exact: the ids
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


BLOCK_FLOATS_PER_STREAM = 68      # include/awd_b200.h AWD_RACE_BLOCK_FLOATS_PER_STREAM


class VocabShardedGenerator:
    """Generation with the decoder rows split over the ranks of `group` (or over `logical_shards` slices on one GPU).

    Per token every rank enqueues: decoder + race over its own rows -> its block of the exchange buffer; ONE in-place
    all-gather of the blocks (68 floats per stream and rank: id, key, keys of the disallowed ids it owns); the merge, which
    gives every rank the same id; one LSTM step on that id (the recurrence is replicated).  Nothing returns to the host
    between tokens -- no clone, no synchronisation -- so the stream runs at the GPU's pace; with `use_graph` the whole
    n_new-token sequence (NCCL included) is captured once into a CUDA graph and replayed per call, which removes the host's
    ~10 launches per token as well."""

    def __init__(self, engine: Engine, bs: int, group=None, logical_shards: Optional[int] = None, use_graph: bool = False):
        import torch.distributed as dist
        self.e, self.bs, self.group, self.logical = engine, bs, group, logical_shards
        if logical_shards is None:
            self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
            self.my = [(self.rank, shard_bounds(engine.vocab_sz, self.world)[self.rank])]
        else:
            self.world, self.rank = logical_shards, 0
            self.my = list(enumerate(shard_bounds(engine.vocab_sz, self.world)))
        self.blocks = torch.zeros(self.world, bs * BLOCK_FLOATS_PER_STREAM, dtype=torch.float32, device=engine.device)
        self.use_graph = use_graph
        self._graphs = {}

    def _steps(self, n_new: int, s, out: torch.Tensor):
        import torch.distributed as dist
        e, lib, st = self.e, self.e.lib, self.e._stream()
        ptr = lambda t: C.c_void_p(t.data_ptr())
        for i in range(n_new):
            for slot, (v0, v1) in self.my:
                _lib.check(lib.awd_decode_race_shard_packed(e._h, v0, v1, C.byref(s), i, ptr(self.blocks[slot]), st))
            if self.logical is None and self.world > 1:
                dist.all_gather_into_tensor(self.blocks.view(-1), self.blocks[self.rank], group=self.group)     # in place
            _lib.check(lib.awd_merge_race_packed(ptr(self.blocks), self.world, self.bs, e.vocab_sz, C.byref(s), ptr(out[i]), st))
            e.forward(out[i:i + 1], need_logits=False, need_raw=False)   # feed back (main.py:71), replicated

    def generate(self, prompt: torch.Tensor, n_new: int, sampling: SamplingParams, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """prompt int64 (P,bs) on the engine's GPU, identical on every rank -> ids (n_new,bs), identical on every rank."""
        e = self.e
        P, bs = prompt.shape
        assert bs == self.bs
        if out is None:
            out = torch.empty(n_new, bs, dtype=torch.int64, device=e.device)
        s, keep = sampling.to_c()
        e.reset(bs)
        e.forward(prompt, need_logits=False, need_raw=False)          # prompt, replicated
        if not self.use_graph:
            self._steps(n_new, s, out)
            return out
        # CUDA graph: the per-token kernels carry their step index and the engine's ping-pong parity as launch arguments, so
        # one graph holds all n_new tokens; it is valid for calls that start from the same state parity (reset + P prompt steps)
        # and the same sampling parameters.
        key = (n_new, P & 1, sampling.seed, sampling.greedy, sampling.temperature, sampling.n_draws, tuple(sampling.banned), out.data_ptr())
        if key not in self._graphs:
            self._steps(1, s, out[:1].clone())                        # warm-up outside capture: workspaces, NCCL channels
            e.reset(bs)
            e.forward(prompt, need_logits=False, need_raw=False)
            torch.cuda.synchronize(e.device)
            g = torch.cuda.CUDAGraph()
            # (thread_local: the NCCL watchdog thread of torch.distributed polls CUDA events while this thread captures)
            with torch.cuda.graph(g, capture_error_mode='thread_local'):
                self._steps(n_new, s, out)
            self._graphs[key] = (g, keep)
            e.reset(bs)                                               # capture ran nothing: redo the prompt and replay
            e.forward(prompt, need_logits=False, need_raw=False)
        self._graphs[key][0].replay()
        # (the engine's host-side bookkeeping was advanced by the capture, not by the replay: reset the engine before using it for
        #  anything but another generate() of this class)
        return out


def generate_vocab_sharded(engine: Engine, prompt: torch.Tensor, n_new: int, sampling: SamplingParams,
                           group=None, logical_shards: Optional[int] = None) -> torch.Tensor:
    """prompt int64 (P,bs), identical on every rank -> ids (n_new,bs), identical on every rank."""
    gen = VocabShardedGenerator(engine, prompt.shape[1], group=group, logical_shards=logical_shards)
    return gen.generate(prompt.to(engine.device), n_new, sampling)
