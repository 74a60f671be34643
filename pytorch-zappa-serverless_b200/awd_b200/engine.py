"""Engine: a Python handle on one awd_engine (include/awd_b200.h).

PyTorch is used here only as plumbing: it owns the output tensors, names the CUDA stream and
moves bytes.  All arithmetic happens in libawd_b200.so.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Mapping, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib


@dataclass
class SamplingParams:
    """The sampling rule of sample_model (reference main.py:61-68); defaults are the reference's."""
    temperature: float = 1.0
    greedy: bool = False
    n_draws: int = 10
    ban_zero: bool = True
    banned: Sequence[int] = field(default_factory=tuple)
    seed: int = 0

    def to_c(self):
        arr = (C.c_int64 * max(len(self.banned), 1))(*[int(b) for b in self.banned])
        s = _lib.AwdSampling(float(self.temperature), int(self.greedy), int(self.n_draws), int(self.ban_zero),
                             len(self.banned), 0, C.cast(arr, C.POINTER(C.c_int64)), int(self.seed) & (2 ** 64 - 1))
        return s, arr   # keep arr alive while s is in use


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


class Engine:
    def __init__(self, vocab_sz: int, emb_sz: int, n_hid: int, n_layers: int, tie_weights: bool = True,
                 bias: bool = True, precision: str = 'fp32', device: int = 0):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise RuntimeError('awd_b200 needs a CUDA device (sm_100a); there is no CPU fallback')
        self.vocab_sz, self.emb_sz, self.n_hid, self.n_layers = vocab_sz, emb_sz, n_hid, n_layers
        self.tie_weights, self.bias = bool(tie_weights), bool(bias)
        self.precision = precision
        self.device = torch.device('cuda', device)
        cfg = _lib.AwdConfig(vocab_sz, emb_sz, n_hid, n_layers, int(tie_weights), int(bias),
                             {'fp32': _lib.PREC_FP32, 'bf16': _lib.PREC_BF16}[precision], device)
        h = C.c_void_p()
        _lib.check(self.lib.awd_create(C.byref(cfg), C.byref(h)))
        self._h = h
        self.layer_hidden = [n_hid if l != n_layers - 1 else emb_sz for l in range(n_layers)]

    def close(self):
        if getattr(self, '_h', None):
            self.lib.awd_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ weights
    def load_state_dict(self, sd: Mapping[str, object]):
        """Strict load of a reference state_dict: unexpected keys raise
        ValueError, missing keys raise at commit."""
        for k, v in sd.items():
            if isinstance(v, torch.Tensor):
                t = v.detach().to(torch.float32).contiguous()
                ptr, shape = t.data_ptr(), tuple(t.shape)
            else:
                t = np.ascontiguousarray(v, dtype=np.float32)
                ptr, shape = t.ctypes.data, t.shape
            shp = (C.c_int64 * len(shape))(*shape)
            _lib.check(self.lib.awd_load_param(self._h, k.encode(), C.c_void_p(ptr), shp, len(shape)))
        _lib.check(self.lib.awd_commit_weights(self._h))

    # ------------------------------------------------------------------ state
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @property
    def bs(self) -> int:
        return int(self.lib.awd_current_bs(self._h))

    def reset(self, bs: int):
        _lib.check(self.lib.awd_reset(self._h, int(bs), self._stream()))

    def get_hidden(self) -> List[Tuple[torch.Tensor, torch.Tensor]]:
        out = []
        bs = self.bs
        for l, H in enumerate(self.layer_hidden):
            h = torch.empty(1, bs, H, dtype=torch.float32, device=self.device)
            c = torch.empty(1, bs, H, dtype=torch.float32, device=self.device)
            _lib.check(self.lib.awd_get_state(self._h, l, _ptr(h), _ptr(c), self._stream()))
            out.append((h, c))
        return out

    def set_hidden(self, hidden: Sequence[Tuple[torch.Tensor, torch.Tensor]]):
        bs = self.bs
        for l, (h, c) in enumerate(hidden):
            H = self.layer_hidden[l]
            if tuple(h.shape) != (1, bs, H) or tuple(c.shape) != (1, bs, H):
                raise ValueError(f'hidden[{l}] must be two tensors of shape (1,{bs},{H})')
            h = h.detach().to(torch.float32).contiguous()
            c = c.detach().to(torch.float32).contiguous()
            _lib.check(self.lib.awd_set_state(self._h, l, _ptr(h), _ptr(c), self._stream()))
            if not h.is_cuda or not c.is_cuda:
                torch.cuda.current_stream(self.device).synchronize()

    # ------------------------------------------------------------------ compute
    def gather(self, tokens: torch.Tensor) -> torch.Tensor:
        tok = tokens.detach().to(torch.int64).contiguous()
        out = torch.empty(*tok.shape, self.emb_sz, dtype=torch.float32, device=self.device)
        _lib.check(self.lib.awd_gather(self._h, _ptr(tok), tok.numel(), _ptr(out), self._stream()))
        if not tok.is_cuda:
            torch.cuda.current_stream(self.device).synchronize()   # the host buffer may go away
        return out

    def forward(self, tokens: torch.Tensor, need_logits: bool = True, need_raw: bool = True,
                last_step_only: bool = False):
        """tokens int64 (sl,bs) on CPU or on the engine's GPU -> (logits or None, raw_outputs or None),
        device tensors.  Carries / updates the hidden state like RNNCore.forward."""
        if tokens.dim() != 2:
            raise ValueError('input must be a LongTensor of shape (sl, bs)')
        tok = tokens.detach().to(torch.int64).contiguous()
        sl, bs = tok.shape
        logits = None
        if need_logits:
            rows = bs if last_step_only else sl * bs
            logits = torch.empty(rows, self.vocab_sz, dtype=torch.float32, device=self.device)
        raws, raw_ptrs = None, None
        if need_raw:
            raws = [torch.empty(sl, bs, H, dtype=torch.float32, device=self.device) for H in self.layer_hidden]
            raw_ptrs = (C.c_void_p * self.n_layers)(*[r.data_ptr() for r in raws])
        _lib.check(self.lib.awd_forward(self._h, _ptr(tok), sl, bs, _ptr(logits), raw_ptrs, int(last_step_only),
                                        self._stream()))
        if not tok.is_cuda:
            # the ids were staged from pageable memory, which already synchronised; range errors
            # are cheap to surface here, as torch's F.embedding would
            _lib.check(self.lib.awd_check_device_errors(self._h, self._stream()))
        return logits, raws

    def decode(self, x: torch.Tensor) -> torch.Tensor:
        """LinearDecoder alone: x (rows, emb_sz) -> logits (rows, vocab_sz)."""
        x = x.detach().to(self.device, torch.float32).contiguous()
        if x.dim() != 2 or x.shape[1] != self.emb_sz:
            raise ValueError('decoder input must be (rows, emb_sz)')
        out = torch.empty(x.shape[0], self.vocab_sz, dtype=torch.float32, device=self.device)
        _lib.check(self.lib.awd_decode(self._h, _ptr(x), x.shape[0], _ptr(out), self._stream()))
        return out

    def check_errors(self):
        _lib.check(self.lib.awd_check_device_errors(self._h, self._stream()))

    def sample(self, logits: torch.Tensor, sampling: SamplingParams, step: int = 0) -> torch.Tensor:
        z = logits.detach().to(self.device, torch.float32).contiguous()
        if z.dim() != 2 or z.shape[1] != self.vocab_sz:
            raise ValueError('logits must be (bs, vocab_sz)')
        out = torch.empty(z.shape[0], dtype=torch.int64, device=self.device)
        s, keep = sampling.to_c()
        _lib.check(self.lib.awd_sample(self._h, _ptr(z), z.shape[0], C.byref(s), int(step), _ptr(out), self._stream()))
        return out

    def generate(self, prompt: torch.Tensor, n_new: int, sampling: SamplingParams,
                 out: Optional[torch.Tensor] = None, sync: bool = True,
                 prompt_len: Optional[Sequence[int]] = None) -> torch.Tensor:
        """prompt int64 (P,bs) -> generated ids (n_new,bs).  `prompt` / `out` may live in (pinned)
        host memory or on the GPU; the library copies them itself (awd_generate)."""
        if prompt.dim() != 2:
            raise ValueError('prompt must be (P, bs)')
        tok = prompt.detach().to(torch.int64).contiguous()
        P, bs = tok.shape
        if out is None:
            out = torch.empty(n_new, bs, dtype=torch.int64, device=self.device)
        elif tuple(out.shape) != (n_new, bs) or out.dtype != torch.int64 or not out.is_contiguous():
            raise ValueError('out must be a contiguous int64 (n_new, bs) tensor')
        s, keep = sampling.to_c()
        if prompt_len is not None:       # ragged: prompts right-aligned in `prompt`, stream b has prompt_len[b] tokens
            if len(prompt_len) != bs:
                raise ValueError('prompt_len must have one entry per stream')
            lens = (C.c_int32 * bs)(*[int(x) for x in prompt_len])
            _lib.check(self.lib.awd_generate_ragged(self._h, _ptr(tok), lens, P, bs, int(n_new), C.byref(s), _ptr(out),
                                                    self._stream()))
        else:
            _lib.check(self.lib.awd_generate(self._h, _ptr(tok), P, bs, int(n_new), C.byref(s), _ptr(out), self._stream()))
        if sync or not tok.is_cuda or not out.is_cuda:
            _lib.check(self.lib.awd_check_device_errors(self._h, self._stream()))
        return out

    def generate_resume(self, bs: int, n_new: int, sampling: SamplingParams, step_base: int,
                        reset: Optional[Sequence[bool]] = None, forced: Optional[torch.Tensor] = None,
                        forced_len: Optional[Sequence[int]] = None, out: Optional[torch.Tensor] = None,
                        sync: bool = True) -> torch.Tensor:
        """Continuous batching (awd_generate_resume): n_new more {decode, draw, feed} steps for the bs slots, whose state
        the engine keeps between calls.  reset[b] zeroes slot b first (a newly admitted request); forced int64 (F,bs) +
        forced_len: prompt tokens fed instead of the drawn ids while a slot is inside its prompt.  -> ids (n_new,bs)."""
        if out is None:
            out = torch.empty(n_new, bs, dtype=torch.int64, device=self.device)
        elif tuple(out.shape) != (n_new, bs) or out.dtype != torch.int64 or not out.is_contiguous():
            raise ValueError('out must be a contiguous int64 (n_new, bs) tensor')
        mask = None
        if reset is not None:
            if len(reset) != bs:
                raise ValueError('reset must have one entry per slot')
            import numpy as np
            mask_np = np.ascontiguousarray(np.asarray(reset, dtype=np.uint8))       # kept alive until the call returns
            mask = mask_np.ctypes.data_as(C.POINTER(C.c_uint8))
        F, fptr, flen = 0, None, None
        if forced is not None:
            forced = forced.detach().to(torch.int64).contiguous()
            if forced.dim() != 2 or forced.shape[1] != bs or forced_len is None or len(forced_len) != bs:
                raise ValueError('forced must be (F, bs) with one forced_len per slot')
            F, fptr = int(forced.shape[0]), _ptr(forced)
            import numpy as np
            flen_np = np.ascontiguousarray(np.asarray(forced_len, dtype=np.int32))
            flen = flen_np.ctypes.data_as(C.POINTER(C.c_int32))
        s, keep = sampling.to_c()
        _lib.check(self.lib.awd_generate_resume(self._h, mask, fptr, flen, F, int(bs), int(n_new), C.byref(s),
                                                int(step_base), _ptr(out), self._stream()))
        if sync:          # sync=False: nothing waits for the stream (host arguments were copied into pinned staging)
            _lib.check(self.lib.awd_check_device_errors(self._h, self._stream()))
        return out

    PROF_CATEGORIES = ('xproj', 'step', 'decoder', 'sample', 'layer0', 'layer1', 'layer2plus')

    def profile(self, enable: bool):
        _lib.check(self.lib.awd_profile(self._h, int(enable)))

    def profile_read(self) -> Dict[str, Tuple[float, int]]:
        """category -> (total ms, launches) of the recording started by profile(True)."""
        out = {}
        for i, name in enumerate(self.PROF_CATEGORIES):
            ms, n = C.c_double(), C.c_int64()
            _lib.check(self.lib.awd_profile_read(self._h, i, C.byref(ms), C.byref(n)))
            out[name] = (ms.value, n.value)
        return out

    def launch_count(self) -> int:
        return int(self.lib.awd_launch_count())
