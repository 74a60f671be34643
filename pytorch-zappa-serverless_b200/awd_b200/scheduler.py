"""Batch scheduler behind ``GET /inference``.

The reference serves one request per thread, each building its own model (main.py:105-112 ->
84-103).  Here every GPU holds ONE model replica; WSGI threads enqueue requests and block on a
future, and one worker thread per replica drains the queue into batched device-side generations
(``awd_generate``).  Request streams are independent (no term of RNNCore.forward couples two
batch rows), so replicas share nothing and no collective is involved.

Turning ids into text is a second pipeline stage with its own thread: for a full device batch (1024 streams x
256 tokens) the per-word rules of main.py:70-79 cost about as much host time as the GPU needs for the batch, so
a worker hands (requests, ids) to the formatter and goes straight back to the queue; the formatter uses the
tabulated `Detokenizer` (text.py), ~5 ms per full batch instead of ~28 ms.
"""
from __future__ import annotations

import os
import queue
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import torch

from .engine import SamplingParams
from .text import Detokenizer, exclude_token_ids


@dataclass
class _Request:
    prompt_words: List[str]
    l: int
    greedy: bool
    temperature: float
    seed: Optional[int]
    future: Future = field(default_factory=Future)

    def group_key(self):
        # streams in one device batch share the sampling rule (prompt lengths may differ: ragged prompts);
        # a request that pins its seed runs alone so that its result does not depend on its neighbours
        return (self.greedy, self.temperature, id(self) if self.seed is not None else None)


class BatchScheduler:
    def __init__(self, models, max_batch: int = 1024, max_wait_ms: float = 2.0):
        self.models = list(models) if isinstance(models, (list, tuple)) else [models]
        self.max_batch, self.max_wait = max_batch, max_wait_ms / 1e3
        self.q: 'queue.Queue[_Request]' = queue.Queue()
        self._stop = False
        self.stats = {'batches': 0, 'requests': 0, 'tokens': 0, 'busy_s': 0.0, 'format_s': 0.0}
        self._fmt_q: 'queue.Queue' = queue.Queue()
        self._detok = {}                     # id(itos) -> Detokenizer (replicas share one vocabulary)
        self._fmt_thread = threading.Thread(target=self._formatter, daemon=True)
        self._fmt_thread.start()
        self._threads = [threading.Thread(target=self._worker, args=(m,), daemon=True) for m in self.models]
        for t in self._threads:
            t.start()

    # ------------------------------------------------------------------ client side
    def submit_async(self, prompt_words: Sequence[str], l: int, greedy: bool = False, temperature: float = 1.0,
                     seed: Optional[int] = None) -> Future:
        r = _Request(list(prompt_words), int(l), bool(greedy), float(temperature), seed)
        self.q.put(r)
        return r.future

    def submit(self, prompt_words: Sequence[str], l: int, **kw) -> str:
        return self.submit_async(prompt_words, l, **kw).result()

    def close(self):
        self._stop = True
        for _ in self._threads:
            self.q.put(None)
        for t in self._threads:
            t.join(timeout=5)
        self._fmt_q.put(None)
        self._fmt_thread.join(timeout=5)

    # ------------------------------------------------------------------ worker side
    def _collect(self) -> List[_Request]:
        first = self.q.get()
        if first is None:
            return []
        batch, key, spill = [first], first.group_key(), []
        deadline = time.monotonic() + self.max_wait
        while len(batch) < self.max_batch:
            try:
                r = self.q.get_nowait()              # the common case under load: no timed wait per request
            except queue.Empty:
                try:
                    r = self.q.get(timeout=max(0.0, deadline - time.monotonic()))
                except queue.Empty:
                    break
            if r is None:
                self.q.put(None)
                break
            (batch if r.group_key() == key else spill).append(r)
        for r in spill:
            self.q.put(r)
        return batch

    def _run_batch(self, model, batch: List[_Request]):
        stoi, itos = model.stoi, model.itos
        n_new = max(r.l for r in batch)
        ids = [[stoi[s] if s in stoi else 0 for s in r.prompt_words] for r in batch]     # main.py:55
        lens = [len(x) for x in ids]
        P = max(lens)
        ids = [[0] * (P - len(x)) + x for x in ids]                                      # right-aligned
        prompt = torch.tensor(ids, dtype=torch.int64).t().contiguous().view(P, len(batch))
        seed = batch[0].seed if batch[0].seed is not None else int.from_bytes(os.urandom(8), 'little')
        sp = SamplingParams(temperature=batch[0].temperature, greedy=batch[0].greedy, n_draws=10, ban_zero=True,
                            banned=exclude_token_ids(stoi), seed=seed)
        ragged = min(lens) != P
        out = (model.generate_ids(prompt, n_new, sp, prompt_len=lens) if ragged else model.generate_ids(prompt, n_new, sp)).cpu()
        self._fmt_q.put((model, batch, out))          # text is built off this thread; the GPU gets its next batch meanwhile
        return n_new * len(batch)

    def _formatter(self):
        while True:
            item = self._fmt_q.get()
            if item is None:
                break
            model, batch, out = item
            t0 = time.monotonic()
            try:
                dt = self._detok.get(id(model.itos))
                if dt is None:
                    dt = self._detok[id(model.itos)] = Detokenizer(model.itos)
                texts = dt.format_batch([r.prompt_words for r in batch], out.numpy(), [r.l for r in batch])
                for r, text in zip(batch, texts):
                    r.future.set_result(text)
            except Exception as e:
                for r in batch:
                    if not r.future.done():
                        r.future.set_exception(e)
            self.stats['format_s'] += time.monotonic() - t0

    def _worker(self, model):
        dev = next(model.parameters()).device
        if dev.type == 'cuda':
            torch.cuda.set_device(dev)
        while not self._stop:
            batch = self._collect()
            if not batch:
                break
            t0 = time.monotonic()
            try:
                ntok = self._run_batch(model, batch)
                self.stats['batches'] += 1
                self.stats['requests'] += len(batch)
                self.stats['tokens'] += ntok
            except Exception as e:          # fail the requests of this batch, not the process
                for r in batch:
                    if not r.future.done():
                        r.future.set_exception(e)
            self.stats['busy_s'] += time.monotonic() - t0
