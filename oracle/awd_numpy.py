"""numpy restatement of the reference AWD-LSTM hot path (TEST INFRASTRUCTURE ONLY).

Every function cites the reference lines it follows (paths are inside
``/root/reference``).  The arithmetic is third-party there (PyTorch); this file
restates the published operator definitions in plain numpy so that it runs on
any box, in float32 (the reference's dtype) or float64 (tie-break ground
truth).  Pinned by ``tests/test_oracle_golden.py`` against fixtures produced
from the live reference module by ``oracle/make_golden.py``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

# main.py:45-46
NO_SPACE_WORDS = ["'s", "'ll", ",", "?", ".", "'t", "'m", "n't", "!", "'", "'ve", ";", "http", ":", "/", "\\"]
CAPITALIZE_WORDS = ['.', '!', '\n']
# main.py:47
EXCLUDE_NAMES = ["xxup", "xxfld", "xxrep"]


@dataclass(frozen=True)
class LMConfig:
    """Arguments of get_language_model that shape the arithmetic (pytorch_models/awd_lstm.py:7-14)."""
    vocab_sz: int
    emb_sz: int
    n_hid: int
    n_layers: int
    tie_weights: bool = True
    bias: bool = True

    def layer_dims(self) -> List[Tuple[int, int]]:
        """(input, hidden) per layer: awd_lstm.py:69 -- last layer's hidden is emb_sz."""
        dims = []
        for l in range(self.n_layers):
            i = self.emb_sz if l == 0 else self.n_hid
            h = self.n_hid if l != self.n_layers - 1 else self.emb_sz
            dims.append((i, h))
        return dims


def state_dict_keys(cfg: LMConfig) -> List[str]:
    """Key order of ``get_language_model(...).state_dict()`` (awd_lstm.py:59-60,69-71,149,36)."""
    keys = ["0.encoder.weight", "0.encoder_dp.emb.weight"]
    for l in range(cfg.n_layers):
        p = f"0.rnns.{l}."
        keys += [p + "weight_hh_l0_raw", p + "module.weight_ih_l0", p + "module.weight_hh_l0",
                 p + "module.bias_ih_l0", p + "module.bias_hh_l0"]
    keys.append("1.decoder.weight")
    if cfg.bias:
        keys.append("1.decoder.bias")
    return keys


def state_dict_shapes(cfg: LMConfig) -> Dict[str, Tuple[int, ...]]:
    shp = {"0.encoder.weight": (cfg.vocab_sz, cfg.emb_sz), "0.encoder_dp.emb.weight": (cfg.vocab_sz, cfg.emb_sz)}
    for l, (i, h) in enumerate(cfg.layer_dims()):
        p = f"0.rnns.{l}."
        shp[p + "weight_hh_l0_raw"] = (4 * h, h)
        shp[p + "module.weight_ih_l0"] = (4 * h, i)
        shp[p + "module.weight_hh_l0"] = (4 * h, h)
        shp[p + "module.bias_ih_l0"] = (4 * h,)
        shp[p + "module.bias_hh_l0"] = (4 * h,)
    shp["1.decoder.weight"] = (cfg.vocab_sz, cfg.emb_sz)
    if cfg.bias:
        shp["1.decoder.bias"] = (cfg.vocab_sz,)
    return shp


def init_state_dict(cfg: LMConfig, seed: int = 0, decoder_bias_scale: float = 0.0) -> Dict[str, np.ndarray]:
    """Random-init with the reference constructor's distributions (not its RNG stream):
    encoder/decoder U(-0.1,0.1) (awd_lstm.py:37,73), nn.LSTM default U(-1/sqrt(H),1/sqrt(H)),
    decoder bias 0 (awd_lstm.py:39) unless ``decoder_bias_scale`` asks for a nonzero one (tests)."""
    rng = np.random.default_rng(seed)
    sd: Dict[str, np.ndarray] = {}
    enc = rng.uniform(-0.1, 0.1, (cfg.vocab_sz, cfg.emb_sz)).astype(np.float32)
    sd["0.encoder.weight"] = enc
    sd["0.encoder_dp.emb.weight"] = enc
    for l, (i, h) in enumerate(cfg.layer_dims()):
        k = 1.0 / np.sqrt(h)
        p = f"0.rnns.{l}."
        whh = rng.uniform(-k, k, (4 * h, h)).astype(np.float32)
        sd[p + "weight_hh_l0_raw"] = whh
        sd[p + "module.weight_ih_l0"] = rng.uniform(-k, k, (4 * h, i)).astype(np.float32)
        sd[p + "module.weight_hh_l0"] = whh
        sd[p + "module.bias_ih_l0"] = rng.uniform(-k, k, (4 * h,)).astype(np.float32)
        sd[p + "module.bias_hh_l0"] = rng.uniform(-k, k, (4 * h,)).astype(np.float32)
    sd["1.decoder.weight"] = enc if cfg.tie_weights else rng.uniform(-0.1, 0.1, (cfg.vocab_sz, cfg.emb_sz)).astype(np.float32)
    if cfg.bias:
        sd["1.decoder.bias"] = (decoder_bias_scale * rng.standard_normal(cfg.vocab_sz)).astype(np.float32)
    return sd


@dataclass
class EffectiveParams:
    encoder: np.ndarray                 # (V,E)
    w_ih: List[np.ndarray]              # (4H_l, I_l)
    w_hh: List[np.ndarray]              # (4H_l, H_l)
    b_ih: List[np.ndarray]
    b_hh: List[np.ndarray]
    decoder_w: np.ndarray               # (V,E)
    decoder_b: Optional[np.ndarray]     # (V,) or None


def effective_params(sd: Dict[str, np.ndarray], cfg: LMConfig, dtype=np.float32) -> EffectiveParams:
    """Which tensors the forward pass really uses after ``load_state_dict`` (main.py:99).

    ``weight_hh_l0_raw`` shares storage with ``module.weight_hh_l0`` (awd_lstm.py:149 builds
    it from ``w.data``) and load copies keys in order, so ``module.weight_hh_l0`` (later key)
    is the surviving value.  With ``tie_weights`` the encoder, ``encoder_dp.emb`` and decoder
    weights are one Parameter (awd_lstm.py:40,59-60) and ``1.decoder.weight`` is copied last.
    """
    a = lambda k: np.asarray(sd[k], dtype=dtype)
    if cfg.tie_weights:
        enc = dec = a("1.decoder.weight")
    else:
        # encoder and encoder_dp.emb are one Parameter; the later key wins
        enc = a("0.encoder_dp.emb.weight")
        dec = a("1.decoder.weight")
    w_ih, w_hh, b_ih, b_hh = [], [], [], []
    for l in range(cfg.n_layers):
        p = f"0.rnns.{l}.module."
        w_ih.append(a(p + "weight_ih_l0"))
        w_hh.append(a(p + "weight_hh_l0"))
        b_ih.append(a(p + "bias_ih_l0"))
        b_hh.append(a(p + "bias_hh_l0"))
    return EffectiveParams(enc, w_ih, w_hh, b_ih, b_hh, dec, a("1.decoder.bias") if cfg.bias else None)


def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def zero_hidden(cfg: LMConfig, bs: int, dtype=np.float32):
    """RNNCore.reset / _one_hidden (awd_lstm.py:93-103): (h,c) zeros of shape (1,bs,H_l) per layer."""
    return [(np.zeros((1, bs, h), dtype), np.zeros((1, bs, h), dtype)) for _, h in cfg.layer_dims()]


def embedding(tokens: np.ndarray, table: np.ndarray) -> np.ndarray:
    """EmbeddingDropout.forward in eval mode (awd_lstm.py:114-122): plain row gather; the
    padding_idx argument has no forward effect."""
    return table[tokens]


def lstm_layer(x: np.ndarray, h0: np.ndarray, c0: np.ndarray, w_ih, w_hh, b_ih, b_hh):
    """One single-layer nn.LSTM call as made at awd_lstm.py:162 (module built at :69-70).

    x (sl,bs,I); h0,c0 (1,bs,H) -> y (sl,bs,H), (hT,cT) each (1,bs,H).
    gates = x_t W_ih^T + b_ih + h_{t-1} W_hh^T + b_hh, chunked i,f,g,o;
    c_t = sig(f) c_{t-1} + sig(i) tanh(g);  h_t = sig(o) tanh(c_t).
    """
    sl, bs, _ = x.shape
    H = w_hh.shape[1]
    h, c = h0[0].copy(), c0[0].copy()
    xp = x.reshape(sl * bs, -1) @ w_ih.T + (b_ih + b_hh)
    xp = xp.reshape(sl, bs, 4 * H)
    y = np.empty((sl, bs, H), dtype=x.dtype)
    for t in range(sl):
        g = xp[t] + h @ w_hh.T
        i_g, f_g, g_g, o_g = g[:, :H], g[:, H:2 * H], g[:, 2 * H:3 * H], g[:, 3 * H:]
        c = _sigmoid(f_g) * c + _sigmoid(i_g) * np.tanh(g_g)
        h = _sigmoid(o_g) * np.tanh(c)
        y[t] = h
    return y, (h[None], c[None])


def rnn_core_forward(p: EffectiveParams, cfg: LMConfig, tokens: np.ndarray, hidden):
    """RNNCore.forward in eval mode (awd_lstm.py:77-91): gather, then n_layers LSTM calls with the
    carried state; dropouts are identity (awd_lstm.py:119,131-132).  Returns raw_outputs and the
    new hidden list (outputs == raw_outputs in eval)."""
    raw = embedding(tokens, p.encoder)
    raw_outputs, new_hidden = [], []
    for l in range(cfg.n_layers):
        raw, hc = lstm_layer(raw, hidden[l][0], hidden[l][1], p.w_ih[l], p.w_hh[l], p.b_ih[l], p.b_hh[l])
        raw_outputs.append(raw)
        new_hidden.append(hc)
    return raw_outputs, new_hidden


def decoder(p: EffectiveParams, out_last: np.ndarray) -> np.ndarray:
    """LinearDecoder.forward (awd_lstm.py:42-46): raw logits, row index = t*bs + b."""
    sl, bs, E = out_last.shape
    z = out_last.reshape(sl * bs, E) @ p.decoder_w.T
    if p.decoder_b is not None:
        z = z + p.decoder_b
    return z


def forward(p: EffectiveParams, cfg: LMConfig, tokens: np.ndarray, hidden):
    """SequentialRNN.__call__ (awd_lstm.py:23-27 -> 77-91 -> 42-46).
    tokens int (sl,bs) -> (decoded (sl*bs,V), raw_outputs, new_hidden)."""
    raw_outputs, new_hidden = rnn_core_forward(p, cfg, tokens, hidden)
    return decoder(p, raw_outputs[-1]), raw_outputs, new_hidden


# ---------------------------------------------------------------- generation (main.py:40-81)

def exclude_tokens(stoi: Dict[str, int]) -> List[int]:
    """main.py:47."""
    return [stoi[t] for t in EXCLUDE_NAMES if t in stoi]


def pick_first_allowed(order: Sequence[int], banned: Sequence[int]) -> int:
    """main.py:63-68: first candidate that is not id 0 and not excluded, else the first candidate."""
    for r in order:
        if int(r) != 0 and int(r) not in banned:
            return int(r)
    return int(order[0])


def topk_order(z: np.ndarray, k: int = 10) -> np.ndarray:
    """Greedy stand-in for main.py:61: the k largest logits, descending,
    ties broken toward the lower index."""
    k = min(k, z.shape[0])
    idx = np.lexsort((np.arange(z.shape[0]), -z))
    return idx[:k]


def generate_ids(p: EffectiveParams, cfg: LMConfig, prompt_ids: np.ndarray, n_new: int,
                 banned: Sequence[int], dtype=np.float32, rng: Optional[np.random.Generator] = None):
    """The id-level part of sample_model (main.py:50-71) for a batch of independent streams.

    prompt_ids (P,bs) are fed one step at a time (main.py:54-57); then n_new times: candidates
    = multinomial(exp(logits),10) when ``rng`` is given (main.py:61: an exponential race, i.e.
    the 10 largest of logits - log(Exp(1))), else the greedy top-10; pick_first_allowed; feed back.
    Returns ids (n_new,bs) and the per-step top-1/top-2 allowed-logit gap (greedy only).
    """
    P, bs = prompt_ids.shape
    hidden = zero_hidden(cfg, bs, dtype)
    z = None
    for t in range(P):
        z, _, hidden = forward(p, cfg, prompt_ids[t:t + 1], hidden)
    out = np.zeros((n_new, bs), np.int64)
    gaps = np.zeros((n_new, bs))
    for i in range(n_new):
        for b in range(bs):
            zb = z[b]
            if rng is None:
                order = topk_order(zb, 10)
                allowed = zb.copy()
                allowed[[0] + list(banned)] = -np.inf
                s = np.sort(allowed)
                gaps[i, b] = s[-1] - s[-2]
            else:
                keys = zb.astype(np.float64) - np.log(rng.exponential(size=zb.shape[0]))
                order = topk_order(keys, 10)
            out[i, b] = pick_first_allowed(order, banned)
        z, _, hidden = forward(p, cfg, out[i:i + 1], hidden)
    return out, gaps


def first_allowed_law(z: np.ndarray, banned: Sequence[int]) -> np.ndarray:
    """Exact law of main.py:61-68 when fewer than 10 ids are disallowed: the first allowed
    element of a Plackett-Luce order is one categorical draw over the allowed set
    (SURVEY.md §7 hard part 6)."""
    zz = z.astype(np.float64).copy()
    zz[[0] + list(banned)] = -np.inf
    e = np.exp(zz - zz.max())
    return e / e.sum()


def format_tokens(prompt_words: Sequence[str], words: Sequence[str]) -> str:
    """String assembly of sample_model (main.py:52,57,58,70-79)."""
    final_string = ''
    for s in prompt_words:
        final_string = final_string + ' ' + s
    last_word = None
    for word in words:
        word = word.capitalize() if last_word in CAPITALIZE_WORDS else word
        if word in NO_SPACE_WORDS:
            final_string = final_string + word
        else:
            final_string = final_string + ' ' + word
        last_word = word
    return final_string


def sample_model(p: EffectiveParams, cfg: LMConfig, itos: Sequence[str], input_words: Sequence[str], l: int = 50,
                 dtype=np.float32, rng: Optional[np.random.Generator] = None) -> str:
    """sample_model(model, input_words, l) (main.py:40-81) at batch 1; OOV prompt words map to 0 (main.py:55)."""
    stoi = {w: i for i, w in enumerate(itos)}
    banned = exclude_tokens(stoi)
    prompt = np.array([[stoi.get(s, 0)] for s in input_words], np.int64).reshape(len(input_words), 1)
    ids, _ = generate_ids(p, cfg, prompt, l, banned, dtype, rng)
    return format_tokens(input_words, [itos[int(i)] for i in ids[:, 0]])


def synthetic_itos(vocab_sz: int) -> List[str]:
    """SURVEY.md §8d vocabulary: specials first so the exclusion and id-0 rules are exercised,
    a few punctuation tokens so the detokeniser rules fire, then w{i}."""
    head = ['xxunk', 'xxpad', 'xxbos', 'xxfld', 'xxup', 'xxrep', '.', ',', '!', '\n', "'s", "n't", '?']
    itos = head[:vocab_sz]
    itos += [f'w{i}' for i in range(len(itos), vocab_sz)]
    return itos
