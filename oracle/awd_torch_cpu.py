"""torch-CPU port of the reference hot path (TEST INFRASTRUCTURE / CPU BASELINE ONLY).

The reference's arithmetic is PyTorch's CPU operators called from
``pytorch_models/awd_lstm.py`` and ``main.py``; ``/root/reference`` does not travel to the GPU
box, so this file re-issues the same operator sequence (``F.embedding`` awd_lstm.py:121,
``nn.LSTM`` :69-70/:162, ``nn.Linear`` :36/:45, ``torch.multinomial(res[-1].exp(), 10)``
main.py:61, first-allowed rule main.py:63-68) on a state_dict.  It is what ``bench.py`` times
as ``cpu_baseline`` (kind "port") and under ``--impl reference``.  Checked against the live
reference in ``oracle/make_golden.py`` and against ``awd_numpy`` in ``tests/``.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .awd_numpy import LMConfig, effective_params


class TorchCpuLM:
    """Holds nn.LSTM / nn.Linear modules filled from the *effective* parameters and carries
    the hidden state between calls like RNNCore.hidden (awd_lstm.py:90)."""

    def __init__(self, sd: Dict[str, np.ndarray], cfg: LMConfig):
        p = effective_params(sd, cfg)
        self.cfg = cfg
        self.enc = torch.from_numpy(np.ascontiguousarray(p.encoder))
        self.rnns: List[nn.LSTM] = []
        for l, (i, h) in enumerate(cfg.layer_dims()):
            m = nn.LSTM(i, h, 1)
            with torch.no_grad():
                m.weight_ih_l0.copy_(torch.from_numpy(p.w_ih[l]))
                m.weight_hh_l0.copy_(torch.from_numpy(p.w_hh[l]))
                m.bias_ih_l0.copy_(torch.from_numpy(p.b_ih[l]))
                m.bias_hh_l0.copy_(torch.from_numpy(p.b_hh[l]))
            m.eval()
            self.rnns.append(m)
        self.dec_w = torch.from_numpy(np.ascontiguousarray(p.decoder_w))
        self.dec_b = torch.from_numpy(p.decoder_b) if p.decoder_b is not None else None
        self.bs = 1
        self.reset()

    def reset(self):
        self.hidden = [(torch.zeros(1, self.bs, h), torch.zeros(1, self.bs, h)) for _, h in self.cfg.layer_dims()]

    def forward(self, tokens: torch.Tensor):
        """tokens int64 (sl,bs) -> (decoded (sl*bs,V), raw_outputs)."""
        sl, bs = tokens.shape
        if bs != self.bs:                      # awd_lstm.py:79-81
            self.bs = bs
            self.reset()
        raw = F.embedding(tokens, self.enc)
        raws, new_hidden = [], []
        for l, rnn in enumerate(self.rnns):
            raw, hc = rnn(raw, self.hidden[l])
            raws.append(raw)
            new_hidden.append(hc)
        self.hidden = [(h.detach(), c.detach()) for h, c in new_hidden]
        out = raws[-1]
        decoded = F.linear(out.view(sl * bs, out.size(2)), self.dec_w, self.dec_b)
        return decoded, raws

    def generate_step(self, z: torch.Tensor, banned: Sequence[int], greedy: bool = False) -> torch.Tensor:
        """main.py:61-68 for every row of z (bs,V) -> next ids (bs,)."""
        if greedy:
            cand = torch.topk(z, min(10, z.shape[1]), dim=1).indices
        else:
            cand = torch.multinomial(z.exp(), min(10, z.shape[1]))
        ok = cand != 0
        for b in banned:
            ok &= cand != b
        first = torch.where(ok.any(1), ok.float().argmax(1), torch.zeros(z.shape[0], dtype=torch.long))
        return cand.gather(1, first[:, None])[:, 0]

    def generate(self, prompt: torch.Tensor, n_new: int, banned: Sequence[int], greedy: bool = False) -> torch.Tensor:
        """Batched id-level sample_model (main.py:50-71): prompt (P,bs) one step per forward."""
        self.bs = prompt.shape[1]
        self.reset()
        z = None
        for t in range(prompt.shape[0]):
            z, _ = self.forward(prompt[t:t + 1])
        out = torch.zeros(n_new, prompt.shape[1], dtype=torch.long)
        for i in range(n_new):
            out[i] = self.generate_step(z, banned, greedy)
            z, _ = self.forward(out[i:i + 1])
        return out
