"""Drop-in for the reference's ``pytorch_models/awd_lstm.py`` backed by the B200 engine.

Same import path, public names, constructor signatures, parameter registration (so the 19-key
fastai-v1 ``state_dict`` loads with ``strict=True``), implicit hidden state, ``reset()`` and
``forward(input)`` return shapes as the reference file (cited as ``ref:`` line numbers of
/root/reference/pytorch_models/awd_lstm.py).  The ``nn.Embedding`` / ``nn.LSTM`` / ``nn.Linear``
children are kept as *parameter containers* only, so attribute paths and the RNG stream of the
constructor are those of the reference; none of their forward methods run.  Every number is
computed by the hand-written sm_100a kernels in ``libawd_b200.so`` through ``awd_b200.Engine``.

Scope: inference (``model.eval()``).  Training-mode dropout, ``qrnn=True`` and ``bidir=True``
are outside the engine and raise.  There is no CPU fallback: without a CUDA device ``forward``
raises.
"""
from __future__ import annotations

import os
import weakref
from typing import Collection, List, Optional, Tuple

import torch
import torch.nn as nn
from torch import LongTensor, Tensor

__all__ = ['get_language_model', 'SequentialRNN', 'LinearDecoder', 'RNNCore', 'EmbeddingDropout', 'RNNDropout',
           'WeightDropout', 'dropout_mask', 'to_detach', 'is_listy', 'AWD_LSTM']


def is_listy(x) -> bool:
    return isinstance(x, (tuple, list))


def to_detach(b):
    "Recursively detach lists of tensors in `b` (ref:16-19)."
    if is_listy(b):
        return [to_detach(o) for o in b]
    return b.detach() if isinstance(b, Tensor) else b


def dropout_mask(x: Tensor, sz: Collection[int], p: float):
    "Bernoulli(1-p)/(1-p) mask of x's type and size `sz` (ref:136-138). Training utility; unused in eval."
    return x.new_empty(*sz).bernoulli_(1 - p).div_(1 - p)


def _eval_only(module: nn.Module, what: str):
    if module.training:
        raise NotImplementedError(
            f'{what}: training-mode dropout is outside the B200 inference engine; call model.eval() first '
            '(the reference does so in sample_model, main.py:50)')


class _EngineHost:
    """Owns the awd_b200.Engine of one language model and keeps its packed device weights in step
    with the module's Parameters (re-uploads after load_state_dict / .to() / in-place edits)."""

    def __init__(self, model: 'SequentialRNN'):
        self._model = weakref.ref(model)
        self.engine = None
        self._stamp = None

    def _params_stamp(self, sd):
        return tuple((k, v.data_ptr(), v._version, v.device.type) for k, v in sd.items())

    def get(self):
        model = self._model()
        core, dec = model[0], model[1]
        sd = model.state_dict(keep_vars=True)
        stamp = self._params_stamp(sd)
        if self.engine is None:
            from awd_b200 import Engine   # fails loudly if libawd_b200.so is missing
            dev = next(model.parameters()).device
            index = dev.index if dev.type == 'cuda' and dev.index is not None else int(os.environ.get('AWD_DEVICE', torch.cuda.current_device() if torch.cuda.is_available() else 0))
            self.engine = Engine(core.encoder.num_embeddings, core.emb_sz, core.n_hid, core.n_layers,
                                 tie_weights=dec.decoder.weight is core.encoder.weight,
                                 bias=dec.decoder.bias is not None,
                                 precision=os.environ.get('AWD_PRECISION', getattr(model, 'precision', 'fp32')),
                                 device=index)
            self._stamp = None
        if stamp != self._stamp:
            self.engine.load_state_dict({k: v.detach() for k, v in sd.items()})
            self._stamp = stamp
            if core.bs != self.engine.bs:
                self.engine.reset(core.bs)
        return self.engine


def get_language_model(vocab_sz: int, emb_sz: int, n_hid: int, n_layers: int, pad_token: int, tie_weights: bool = True,
                       qrnn: bool = False, bias: bool = True, bidir: bool = False, output_p: float = 0.4,
                       hidden_p: float = 0.2, input_p: float = 0.6, embed_p: float = 0.1, weight_p: float = 0.5):
    "Create a full AWD-LSTM (ref:7-14); same arguments, same module tree."
    core = RNNCore(vocab_sz, emb_sz, n_hid=n_hid, n_layers=n_layers, pad_token=pad_token, qrnn=qrnn, bidir=bidir,
                   hidden_p=hidden_p, input_p=input_p, embed_p=embed_p, weight_p=weight_p)
    tie = core.encoder if tie_weights else None
    return SequentialRNN(core, LinearDecoder(vocab_sz, emb_sz, output_p, tie_encoder=tie, bias=bias))


class SequentialRNN(nn.Sequential):
    "A sequential module that passes the reset call to its children (ref:23-27)."

    def __init__(self, *args):
        super().__init__(*args)
        if len(self) == 2 and isinstance(self[0], RNNCore) and isinstance(self[1], LinearDecoder):
            host = _EngineHost(self)
            object.__setattr__(self, '_awd_host', host)
            object.__setattr__(self[0], '_awd_host', host)
            object.__setattr__(self[1], '_awd_host', host)
            object.__setattr__(self[0].encoder_dp, '_awd_host', host)

    def reset(self):
        for c in self.children():
            if hasattr(c, 'reset'):
                c.reset()

    def forward(self, input):
        """RNNCore.forward then LinearDecoder.forward as ONE engine call (ref:77-91 + 42-46):
        input LongTensor (sl,bs) -> (decoded (sl*bs,V), raw_outputs, outputs)."""
        host = getattr(self, '_awd_host', None)
        if host is None:
            return super().forward(input)
        core = self[0]
        _eval_only(self, 'SequentialRNN.forward')
        logits, raws = core._run(input, need_logits=True)
        return logits, raws, list(raws)

    # ---- additions (not in the reference): the device-side generation loop of main.py:50-71
    def generate_ids(self, prompt: Tensor, n_new: int, sampling=None, out: Optional[Tensor] = None,
                     prompt_len=None) -> Tensor:
        from awd_b200 import SamplingParams
        eng = self._awd_host.get()
        out = eng.generate(prompt, n_new, sampling or SamplingParams(), out=out, prompt_len=prompt_len)
        self[0].bs = prompt.shape[1]
        return out


    def resume_ids(self, bs: int, n_new: int, sampling=None, step_base: int = 0, reset=None, forced=None,
                   forced_len=None, out: Optional[Tensor] = None, sync: bool = True) -> Tensor:
        "Continuous batching (awd_generate_resume): the bs slots keep their state between calls."
        from awd_b200 import SamplingParams
        eng = self._awd_host.get()
        out = eng.generate_resume(bs, n_new, sampling or SamplingParams(), step_base, reset=reset, forced=forced,
                                  forced_len=forced_len, out=out, sync=sync)
        self[0].bs = int(bs)
        return out


class LinearDecoder(nn.Module):
    "To go on top of a RNNCore module and create a Language Model (ref:29-46)."

    initrange = 0.1

    def __init__(self, n_out: int, n_hid: int, output_p: float, tie_encoder=None, bias: bool = True):
        super().__init__()
        self.decoder = nn.Linear(n_hid, n_out, bias=bias)
        with torch.no_grad():
            self.decoder.weight.uniform_(-self.initrange, self.initrange)
        self.output_dp = RNNDropout(output_p)
        if bias:
            with torch.no_grad():
                self.decoder.bias.zero_()
        if tie_encoder:
            self.decoder.weight = tie_encoder.weight

    def forward(self, input: Tuple[List[Tensor], List[Tensor]]) -> Tuple[Tensor, List[Tensor], List[Tensor]]:
        raw_outputs, outputs = input
        _eval_only(self, 'LinearDecoder.forward')
        host = getattr(self, '_awd_host', None)
        if host is None:
            raise RuntimeError('LinearDecoder runs on the B200 engine of a model built by get_language_model')
        last = self.output_dp(outputs[-1])
        decoded = host.get().decode(last.reshape(last.size(0) * last.size(1), last.size(2)))
        return decoded.to(last.device), raw_outputs, outputs


class RNNCore(nn.Module):
    "AWD-LSTM encoder (ref:48-103): embedding, n_layers weight-dropped LSTMs, implicit hidden state."

    initrange = 0.1

    def __init__(self, vocab_sz: int, emb_sz: int, n_hid: int, n_layers: int, pad_token: int, bidir: bool = False,
                 hidden_p: float = 0.2, input_p: float = 0.6, embed_p: float = 0.1, weight_p: float = 0.5,
                 qrnn: bool = False):
        super().__init__()
        if qrnn:
            raise NotImplementedError('qrnn=True: the reference imports pytorch_models.qrnn, which it does not ship '
                                      '(ref:62-63); the B200 engine implements the LSTM path only')
        if bidir:
            raise NotImplementedError('bidir=True is outside the B200 engine (main.py:96 never sets it)')
        self.bs, self.qrnn, self.ndir = 1, qrnn, 1
        self.emb_sz, self.n_hid, self.n_layers = emb_sz, n_hid, n_layers
        self.encoder = nn.Embedding(vocab_sz, emb_sz, padding_idx=pad_token)
        self.encoder_dp = EmbeddingDropout(self.encoder, embed_p)
        sizes = [(emb_sz if l == 0 else n_hid, n_hid if l != n_layers - 1 else emb_sz) for l in range(n_layers)]
        self.rnns = nn.ModuleList([WeightDropout(nn.LSTM(i, h, 1), weight_p) for i, h in sizes])
        with torch.no_grad():
            self.encoder.weight.uniform_(-self.initrange, self.initrange)   # pad row included (ref:73)
        self.input_dp = RNNDropout(input_p)
        self.hidden_dps = nn.ModuleList([RNNDropout(hidden_p) for _ in range(n_layers)])
        self._hidden_cache = None

    # -- hidden state lives on the GPU inside the engine; `hidden` materialises it on demand
    @property
    def hidden(self):
        if self._hidden_cache is None:
            host = getattr(self, '_awd_host', None)
            if host is None or host.engine is None:
                raise AttributeError("'RNNCore' object has no attribute 'hidden' (call reset() or forward first)")
            dev = self._out_device()
            self._hidden_cache = [(h.to(dev), c.to(dev)) for h, c in host.get().get_hidden()]
        return self._hidden_cache

    @hidden.setter
    def hidden(self, value):
        host = getattr(self, '_awd_host', None)
        if host is None:
            raise RuntimeError('RNNCore runs on the B200 engine of a model built by get_language_model')
        eng = host.get()
        value = [(h, c) for h, c in value]
        if value and value[0][0].size(1) != eng.bs:
            eng.reset(value[0][0].size(1))
        eng.set_hidden(value)
        self._hidden_cache = None

    def _out_device(self):
        return self.encoder.weight.device

    def _run(self, input: LongTensor, need_logits: bool):
        host = getattr(self, '_awd_host', None)
        if host is None:
            raise RuntimeError('RNNCore runs on the B200 engine of a model built by get_language_model')
        if input.dim() != 2:
            raise ValueError(f'input must be (sl, bs); got {tuple(input.shape)}')
        sl, bs = input.size()
        eng = host.get()
        if bs != self.bs:                     # implicit reset on batch-size change (ref:79-81)
            self.bs = bs
            self.reset()
        elif eng.bs != bs:                    # caller changed .bs without reset(): state has no meaning
            eng.reset(bs)
        logits, raws = eng.forward(input, need_logits=need_logits, need_raw=True)
        self._hidden_cache = None
        dev = self._out_device()
        if dev.type != 'cuda':
            raws = [r.to(dev) for r in raws]
            logits = logits.to(dev) if logits is not None else None
        return logits, raws

    def forward(self, input: LongTensor) -> Tuple[List[Tensor], List[Tensor]]:
        _eval_only(self, 'RNNCore.forward')
        _, raws = self._run(input, need_logits=False)
        return raws, list(raws)              # eval: outputs are the raw outputs (dropout is identity)

    def _one_hidden(self, l: int) -> Tensor:
        "Return one zero hidden state (ref:93-96)."
        nh = self.n_hid if l != self.n_layers - 1 else self.emb_sz
        return next(self.parameters()).new_zeros(self.ndir, self.bs, nh)

    def reset(self):
        "Reset the hidden states to zero for `self.bs` streams (ref:98-103)."
        for r in self.rnns:
            r.reset()
        self.weights = next(self.parameters()).data
        host = getattr(self, '_awd_host', None)
        if host is not None and torch.cuda.is_available():
            host.get().reset(self.bs)
            self._hidden_cache = None
        else:
            self._hidden_cache = [(self._one_hidden(l), self._one_hidden(l)) for l in range(self.n_layers)]


class EmbeddingDropout(nn.Module):
    "Embedding lookup with (training-only) row dropout (ref:105-122)."

    def __init__(self, emb: nn.Module, embed_p: float):
        super().__init__()
        self.emb, self.embed_p = emb, embed_p
        self.pad_idx = self.emb.padding_idx
        if self.pad_idx is None:
            self.pad_idx = -1

    def forward(self, words: LongTensor, scale: Optional[float] = None) -> Tensor:
        if self.training and self.embed_p != 0:
            _eval_only(self, 'EmbeddingDropout.forward')
        if scale:
            raise NotImplementedError('scale is never passed on the inference path (ref:120)')
        host = getattr(self, '_awd_host', None)
        if host is None:
            raise RuntimeError('EmbeddingDropout runs on the B200 engine of a model built by get_language_model')
        out = host.get().gather(words)
        host.get().check_errors()
        return out.to(self.emb.weight.device)


class RNNDropout(nn.Module):
    "Dropout that is consistent on the seq_len dimension (ref:124-134); identity in eval."

    def __init__(self, p: float = 0.5):
        super().__init__()
        self.p = p

    def forward(self, x: Tensor) -> Tensor:
        if not self.training or self.p == 0.:
            return x
        _eval_only(self, 'RNNDropout.forward')


class WeightDropout(nn.Module):
    """Parameter holder of one weight-dropped layer (ref:140-168): registers `<name>_raw`, a
    Parameter that shares storage with the wrapped module's weight, so state_dict keys and the
    aliasing seen by load_state_dict are the reference's.  In eval the effective W_hh is raw."""

    def __init__(self, module: nn.Module, weight_p: float, layer_names: Collection[str] = ['weight_hh_l0']):
        super().__init__()
        self.module, self.weight_p, self.layer_names = module, weight_p, layer_names
        for name in self.layer_names:
            w = getattr(self.module, name)
            self.register_parameter(f'{name}_raw', nn.Parameter(w.data))

    def forward(self, *args):
        raise NotImplementedError('a single weight-dropped LSTM layer is not a separate entry point of the B200 '
                                  'engine; call RNNCore.forward / the language model')

    def reset(self):
        if hasattr(self.module, 'reset'):
            self.module.reset()


class AWD_LSTM(nn.Module):
    """Addition (not in the reference, named by BASELINE.json): explicit-hidden wrapper.
    ``forward(input, hidden=None) -> (decoded, hidden)`` sets the carried state first when
    `hidden` is given and returns the new one."""

    def __init__(self, vocab_sz: int, emb_sz: int, n_hid: int, n_layers: int, pad_token: int = 1, **kwargs):
        super().__init__()
        self.lm = get_language_model(vocab_sz, emb_sz, n_hid, n_layers, pad_token, **kwargs)

    def reset(self):
        self.lm.reset()

    def forward(self, input: LongTensor, hidden=None):
        if hidden is not None:
            self.lm[0].bs = input.size(1)
            self.lm[0].hidden = hidden
        decoded, _, _ = self.lm(input)
        return decoded, self.lm[0].hidden
