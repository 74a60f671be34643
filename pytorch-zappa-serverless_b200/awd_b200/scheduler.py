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


def _replica_seed(seed: int, replica: int) -> int:
    """Every replica draws from its own Philox stream (splitmix64 of the replica index folded into the seed): two GPUs
    serving the same prompt in the same slot and round must not return the same sampled text."""
    if replica == 0:
        return seed
    z = (replica * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return (seed ^ z ^ (z >> 31)) & 0xFFFFFFFFFFFFFFFF


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
        self._threads = [threading.Thread(target=self._worker, args=(m, i), daemon=True) for i, m in enumerate(self.models)]
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

    def _worker(self, model, replica=0):
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


class ContinuousBatcher:
    """Admit / retire requests every `chunk` tokens instead of per batch.

    The engine keeps `slots` stream slots alive (awd_generate_resume).  Every round the worker retires the slots
    whose requests are complete, admits waiting requests into free slots (their state is reset and their prompt is fed
    as forced tokens inside the device loop, main.py:54-57), and runs `chunk` more tokens for all slots at once.  A request waits at most one chunk for a free
    slot and occupies it only for its own prompt + l tokens, whatever the other requests ask for.  Streams do not
    interact, so in greedy mode a request's words are exactly those it gets alone.  Several replicas (one model per
    GPU) share the queue, one worker each.  All requests share one
    sampling rule (the reference's, or greedy); per-request seeds are not offered here.
    """

    def __init__(self, model, slots: int = 256, chunk: int = 32, greedy: bool = False, temperature: float = 1.0,
                 seed: Optional[int] = None, idle_wait_ms: float = 50.0):
        self.models = list(model) if isinstance(model, (list, tuple)) else [model]      # one replica per GPU, one worker each
        self.model, self.slots, self.chunk = self.models[0], int(slots), int(chunk)
        self.greedy, self.temperature = bool(greedy), float(temperature)
        self.seed = seed if seed is not None else int.from_bytes(os.urandom(8), 'little')
        self.idle_wait = idle_wait_ms / 1e3
        self.q: 'queue.Queue[_Request]' = queue.Queue()
        self._stop = False
        self.stats = {'rounds': 0, 'requests': 0, 'tokens': 0, 'slot_tokens': 0, 'device_call_s': 0.0, 'host_s': 0.0}
        self._detok = Detokenizer(self.model.itos)
        self._fmt_q: 'queue.Queue' = queue.Queue()
        self._fmt_thread = threading.Thread(target=self._formatter, daemon=True)
        self._fmt_thread.start()
        self._threads = [threading.Thread(target=self._worker, args=(m, i), daemon=True) for i, m in enumerate(self.models)]
        for t in self._threads:
            t.start()

    def _formatter(self):
        while True:
            item = self._fmt_q.get()
            if item is None:
                break
            reqs, ids = item
            try:
                texts = self._detok.format_batch([r.prompt_words for r in reqs], ids, [r.l for r in reqs])
                for r, t in zip(reqs, texts):
                    r.future.set_result(t)
            except Exception as e:
                for r in reqs:
                    if not r.future.done():
                        r.future.set_exception(e)

    def submit_async(self, prompt_words: Sequence[str], l: int) -> Future:
        r = _Request(list(prompt_words), int(l), self.greedy, self.temperature, None)
        if not r.prompt_words:
            r.future.set_exception(ValueError('empty prompt'))
            return r.future
        if r.l <= 0:                                     # nothing to generate: the prompt alone (main.py:52-58)
            r.future.set_result(''.join(' ' + w for w in r.prompt_words))
            return r.future
        self.q.put(r)
        return r.future

    def submit(self, prompt_words: Sequence[str], l: int) -> str:
        return self.submit_async(prompt_words, l).result()

    def close(self):
        self._stop = True
        for _ in self._threads:
            self.q.put(None)
        for t in self._threads:
            t.join(timeout=10)
        self._fmt_q.put(None)
        self._fmt_thread.join(timeout=10)

    def _worker(self, model, replica=0):
        S, C = self.slots, self.chunk
        stoi = model.stoi
        dev = next(model.parameters()).device
        if dev.type == 'cuda':
            torch.cuda.set_device(dev)
        sp = SamplingParams(temperature=self.temperature, greedy=self.greedy, n_draws=10, ban_zero=True,
                            banned=exclude_token_ids(stoi), seed=_replica_seed(self.seed, replica))
        import numpy as np
        req: List[Optional[_Request]] = [None] * S       # request in each slot
        pending: dict = {}                               # slot -> prompt tokens still to be forced (few slots at a time)
        want = np.zeros(S, dtype=np.int64)               # words asked for (0: free slot)
        have = np.zeros(S, dtype=np.int64)               # words generated so far
        buf = np.zeros((S, 64 + C), dtype=np.int64)      # generated ids per slot; widened on demand
        free = list(range(S - 1, -1, -1))
        step_base = 0
        t_post = None
        overlapped = dev.type == 'cuda' and os.environ.get('AWD_CONTINUOUS_OVERLAP', '1') != '0'
        out_bufs = [torch.empty(C, S, dtype=torch.int64) for _ in range(2)]       # host side; pinned when rounds overlap
        if overlapped:
            out_bufs = [t.pin_memory() for t in out_bufs]
        inflight = None                                  # (plan, ids buffer, event) of the round still running
        while not self._stop:
            # ---- admit
            reset = None
            while free:
                try:
                    r = self.q.get(timeout=self.idle_wait) if len(free) == S else self.q.get_nowait()
                except queue.Empty:
                    break
                if r is None:
                    self._stop = True
                    break
                b = free.pop()
                if reset is None:
                    reset = np.zeros(S, dtype=np.uint8)
                req[b], reset[b], want[b], have[b] = r, 1, max(r.l, 0), 0
                pending[b] = [stoi[w] if w in stoi else 0 for w in r.prompt_words]      # main.py:55
                if r.l + C > buf.shape[1]:
                    buf = np.concatenate([buf, np.zeros((S, r.l + C - buf.shape[1]), dtype=np.int64)], axis=1)
            if self._stop:
                break
            if len(free) == S:
                continue
            # ---- one chunk for every slot
            forced = forced_len = None
            if pending:
                F = min(C, max(len(p) for p in pending.values()))
                fnp = np.zeros((F, S), dtype=np.int64)
                forced_len = np.zeros(S, dtype=np.int32)
                if F == 1:                               # the reference's own requests: one-word prompts
                    idx = np.fromiter(pending.keys(), dtype=np.int64, count=len(pending))
                    fnp[0, idx] = np.fromiter((p[0] for p in pending.values()), dtype=np.int64, count=len(pending))
                    forced_len[idx] = 1
                else:
                    for b, p in pending.items():
                        n = min(len(p), F)
                        forced_len[b] = n
                        fnp[:n, b] = p[:n]
                forced = torch.from_numpy(fnp)
            # ---- plan this round's bookkeeping from the counts alone (which slots finish does not depend on the ids),
            #      so that the round can be enqueued before the previous round's ids are read back
            nf = np.zeros(S, dtype=np.int64)                         # leading entries of a slot's chunk that are prompt tokens
            for b, p in list(pending.items()):
                nf[b] = min(len(p), C)
                if len(p) > C:
                    pending[b] = p[C:]
                else:
                    del pending[b]
            live = want > 0
            plain = np.nonzero(live & (nf == 0))[0]
            special = [(int(b), int(nf[b]), int(have[b])) for b in np.nonzero(live & (nf > 0))[0]]
            pos_plain = have[plain].copy()
            have[live] += C - nf[live]
            done = np.nonzero(live & (have >= want))[0]
            plan = (plain, pos_plain, special, done, [req[b] for b in done], want[done].copy())
            for b in done:
                req[b] = None
                free.append(int(b))
            want[done] = 0
            t_call = time.monotonic()
            if t_post is not None:
                self.stats['host_s'] += t_call - t_post        # bookkeeping between two device calls
            out_buf = out_bufs[step_base & 1]
            try:
                model.resume_ids(S, C, sp, step_base=step_base, reset=reset, forced=forced, forced_len=forced_len,
                                 out=out_buf, sync=not overlapped)
                ev = None
                if overlapped:
                    ev = torch.cuda.Event()
                    ev.record(torch.cuda.current_stream(dev))
            except Exception as e:                      # fail what is in flight, keep serving
                for r in [x for x in req if x is not None] + plan[4] + (inflight[0][4] if inflight else []):
                    if not r.future.done():
                        r.future.set_exception(e)
                req = [None] * S
                inflight = None
                pending.clear(); want[:] = 0; have[:] = 0
                free = list(range(S - 1, -1, -1))
                continue
            t_post = time.monotonic()
            self.stats['device_call_s'] += t_post - t_call
            step_base += 1
            self.stats['rounds'] += 1
            self.stats['slot_tokens'] += S * C
            # ---- the ids of the PREVIOUS round (this one keeps running meanwhile), or of this one when not overlapped
            todo = [inflight] if inflight is not None else []
            inflight = (plan, out_buf, ev)
            if not overlapped or (len(free) == S and not pending and self.q.empty()):
                todo.append(inflight)                   # nothing left to enqueue behind it: drain
                inflight = None
            for (pl_plain, pl_pos, pl_special, pl_done, pl_reqs, pl_want), obuf, oev in todo:
                if oev is not None:
                    oev.synchronize()
                cols = obuf.numpy().T                                # (S, C)
                if len(pl_plain):                                    # one scatter for all slots past their prompt
                    buf[pl_plain[:, None], pl_pos[:, None] + np.arange(C)] = cols[pl_plain]
                for b, n_forced, pos in pl_special:
                    buf[b, pos:pos + C - n_forced] = cols[b, n_forced:]
                if len(pl_done):
                    L = int(pl_want.max())
                    ids = np.ascontiguousarray(buf[pl_done, :max(L, 1)].T)
                    self._fmt_q.put((pl_reqs, ids))                  # text is built off this thread
                    self.stats['requests'] += len(pl_reqs)
                    self.stats['tokens'] += int(pl_want.sum())
