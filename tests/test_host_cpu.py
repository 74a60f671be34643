"""CPU-side checks (no GPU): the C-ABI library loads and exports what include/*.h declares, the
drop-in module keeps the reference's construction / state_dict contract, host text rules and the
batch scheduler logic.  No compute call is made."""
import os
import re
import sys
import threading

import numpy as np
import pytest
import torch

from helpers import load_golden
from oracle import awd_numpy as on

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from awd_b200 import _lib
    from awd_b200.build import build_library
    build_library()
    hdr = open(os.path.join(ROOT, 'include', 'awd_b200.h')).read()
    declared = set(re.findall(r'\b(awd_[a-z_0-9]+)\s*\(', hdr))
    assert declared, 'no declarations found'
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name)
    assert b'awd_b200' in lib.awd_version()
    assert lib.awd_launch_count() == 0


def test_no_cpu_fallback_without_gpu():
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from awd_b200 import Engine
    with pytest.raises(RuntimeError, match='no CPU fallback|CUDA'):
        Engine(10, 4, 4, 1)
    from pytorch_models import awd_lstm
    m = awd_lstm.get_language_model(10, 4, 4, 2, pad_token=1).eval()
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 1, dtype=torch.long))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'pytorch-zappa-serverless_b200')
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h', '.cpp')):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, re.M), f
                assert '/root/reference' not in src.replace('/root/reference/pytorch_models/awd_lstm.py', ''), f


def test_constructor_reproduces_reference_init_stream_and_state_dict():
    """torch.manual_seed(s); get_language_model(...) gives bit-identical parameters to the
    reference's constructor (fixture built with the live reference under the same seed)."""
    from pytorch_models import awd_lstm
    cfg, sd, _ = load_golden('fwd_tiny')
    torch.manual_seed(1)
    m = awd_lstm.get_language_model(40, 8, 12, 3, pad_token=1)
    with torch.no_grad():
        m[1].decoder.bias.normal_(0, 0.05)
    msd = m.state_dict()
    assert list(msd.keys()) == on.state_dict_keys(cfg) == list(sd.keys())
    for k in sd:
        assert np.array_equal(msd[k].numpy(), sd[k]), k
    # aliasing: tied weights are one Parameter; raw shares storage with the LSTM weight
    assert m[1].decoder.weight is m[0].encoder.weight
    assert m[0].rnns[0].weight_hh_l0_raw.data_ptr() == m[0].rnns[0].module.weight_hh_l0.data_ptr()
    assert m[0].bs == 1 and m[0].emb_sz == 8 and m[0].n_hid == 12 and m[0].n_layers == 3
    cfg2, sd2, _ = load_golden('fwd_untied_l2')
    m2 = awd_lstm.get_language_model(97, 12, 18, 2, pad_token=1, tie_weights=False)
    assert list(m2.state_dict().keys()) == list(sd2.keys())


def test_strict_load_and_alias_last_wins():
    from pytorch_models import awd_lstm
    cfg, sd, _ = load_golden('fwd_alias')
    m = awd_lstm.get_language_model(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, pad_token=1)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in sd.items()})
    assert np.array_equal(m[0].encoder.weight.detach().numpy(), sd['1.decoder.weight'])
    assert np.array_equal(m[0].rnns[1].weight_hh_l0_raw.detach().numpy(), sd['0.rnns.1.module.weight_hh_l0'])
    bad = {k: torch.from_numpy(v.copy()) for k, v in sd.items() if k != '0.rnns.0.module.weight_hh_l0'}
    with pytest.raises(RuntimeError, match='Missing key'):
        m.load_state_dict(bad)
    with pytest.raises(NotImplementedError):
        awd_lstm.get_language_model(10, 4, 4, 2, pad_token=1, qrnn=True)
    m.reset()
    assert [tuple(h.shape) for h, _ in m[0].hidden] == [(1, 1, cfg.n_hid), (1, 1, cfg.n_hid), (1, 1, cfg.emb_sz)]


def test_product_text_rules_equal_oracle():
    from awd_b200.text import format_tokens, exclude_token_ids, NO_SPACE_WORDS, CAPITALIZE_WORDS
    assert set(NO_SPACE_WORDS) == set(on.NO_SPACE_WORDS) and set(CAPITALIZE_WORDS) == set(on.CAPITALIZE_WORDS)
    rng = np.random.default_rng(0)
    vocab = on.synthetic_itos(60) + ["'ll", ';', 'http', ':', '/', '\\', "'", "'ve", "'t", "'m"]
    for _ in range(50):
        words = [vocab[i] for i in rng.integers(0, len(vocab), 30)]
        prompt = [vocab[i] for i in rng.integers(0, len(vocab), 3)]
        assert format_tokens(prompt, words) == on.format_tokens(prompt, words)
    stoi = {w: i for i, w in enumerate(vocab)}
    assert exclude_token_ids(stoi) == on.exclude_tokens(stoi)
    cfg, sd, d = load_golden('gen_greedy_small')
    itos = [str(s) for s in d['itos']]
    assert format_tokens([str(s) for s in d['prompt_words']], [itos[i] for i in d['ids']]) == str(d['text'])


def test_batch_detokenizer_equals_word_loop():
    """The tabulated batch formatter gives format_tokens' strings: capitalisation after
    phrase ends, the no-space test on the CAPITALISED word ('http' -> ' Http'), per-stream lengths."""
    from awd_b200.text import Detokenizer, format_tokens
    vocab = on.synthetic_itos(60) + ["'ll", ';', 'http', ':', '/', '\\', "'", "'ve", "'t", "'m", 'i', "n't"]
    dt = Detokenizer(vocab)
    rng = np.random.default_rng(3)
    ids = rng.integers(0, len(vocab), (40, 33))
    ids[rng.random(ids.shape) < 0.25] = vocab.index('.')
    lens = rng.integers(0, 41, 33)
    prompts = [[vocab[i] for i in rng.integers(0, len(vocab), b % 4)] for b in range(33)]
    want = [on.format_tokens(prompts[b], [vocab[i] for i in ids[:lens[b], b]]) for b in range(33)]
    assert dt.format_batch(prompts, ids, lens) == want
    assert dt.format_batch(prompts, torch.from_numpy(ids), lens) == want
    assert want == [format_tokens(prompts[b], [vocab[i] for i in ids[:lens[b], b]]) for b in range(33)]
    assert dt.format_batch([['a']], np.zeros((0, 1), dtype=np.int64), [0]) == [' a']
    with pytest.raises(ValueError):
        dt.format_batch(prompts, ids[:, :5], lens)


class _FakeModel:
    """Stands in for the GPU model: 'generates' ids that encode (prompt id, step)."""

    def __init__(self):
        self.itos = on.synthetic_itos(50)
        self.stoi = {w: i for i, w in enumerate(self.itos)}
        self.calls = []

    def parameters(self):
        yield torch.zeros(1)

    def generate_ids(self, prompt, n_new, sampling, prompt_len=None):
        self.calls.append((tuple(prompt.shape), n_new, sampling.greedy, prompt_len))
        return (prompt[-1:, :] + torch.arange(1, n_new + 1)[:, None]) % 50


def test_scheduler_batches_and_routes_results():
    from awd_b200.scheduler import BatchScheduler
    from awd_b200.text import format_tokens
    fm = _FakeModel()
    s = BatchScheduler(fm, max_batch=8, max_wait_ms=200)
    futs = [s.submit_async([f'w{20 + i}'], 3 + (i % 2)) for i in range(6)]
    futs.append(s.submit_async(['w20', 'w21'], 2))          # different prompt length: same batch, ragged
    texts = [f.result(timeout=10) for f in futs]
    s.close()
    for i in range(6):
        ids = [(20 + i + k) % 50 for k in range(1, 4 + (i % 2))]
        assert texts[i] == format_tokens([f'w{20 + i}'], [fm.itos[j] for j in ids])
    assert texts[6].startswith(' w20 w21')
    assert sum(c[0][1] for c in fm.calls) == 7 and len(fm.calls) <= 3
    ragged = [c for c in fm.calls if c[3] is not None]
    assert ragged and all(c[0][0] == 2 and sorted(set(c[3])) == [1, 2] for c in ragged)   # right-aligned, lengths passed
    assert s.stats['requests'] == 7


class _FakeSlotModel:
    """resume_ids semantics on the CPU: slot b's next id is (last fed id * 3 + 1) % 47 + 3; forced ids override."""

    def __init__(self):
        self.itos = on.synthetic_itos(50)
        self.stoi = {w: i for i, w in enumerate(self.itos)}
        self.last = None
        self.calls = 0

    def parameters(self):
        yield torch.zeros(1)

    def resume_ids(self, bs, n_new, sampling, step_base=0, reset=None, forced=None, forced_len=None, out=None, sync=True):
        self.calls += 1
        if self.last is None or len(self.last) != bs:
            self.last = [0] * bs
        for b in range(bs):
            if reset is not None and reset[b]:
                self.last[b] = 0
        out = torch.zeros(n_new, bs, dtype=torch.int64) if out is None else out
        for i in range(n_new):
            for b in range(bs):
                nxt = (self.last[b] * 3 + 1) % 47 + 3
                if forced is not None and i < forced.shape[0] and i < forced_len[b]:
                    nxt = int(forced[i, b])
                out[i, b] = nxt
                self.last[b] = nxt
        return out


def test_continuous_batcher_admits_and_retires_per_chunk():
    from awd_b200.scheduler import ContinuousBatcher
    from awd_b200.text import format_tokens
    fm = _FakeSlotModel()
    cb = ContinuousBatcher(fm, slots=3, chunk=4, greedy=True)
    reqs = [([f'w{20 + i}'] * (1 + i % 6), 1 + (3 * i) % 11) for i in range(11)]       # prompts up to 6 tokens (> chunk)
    futs = [cb.submit_async(w, l) for w, l in reqs]
    texts = [f.result(timeout=20) for f in futs]
    cb.close()
    for (words, l), text in zip(reqs, texts):
        last, ids = 0, []
        for w in words:
            last = fm.stoi[w]
        for _ in range(l):
            last = (last * 3 + 1) % 47 + 3
            ids.append(last)
        assert text == format_tokens(words, [fm.itos[i] for i in ids])
    assert cb.stats['requests'] == 11 and cb.stats['rounds'] > 4          # 11 requests through 3 slots
    bad = ContinuousBatcher(fm, slots=2, chunk=2).submit_async([], 3)
    with pytest.raises(ValueError):
        bad.result(timeout=5)


def test_wsgi_route_contract(monkeypatch):
    """GET /inference -> 200 application/json {"response": {"text": str}} (reference main.py:105-112)."""
    import json
    import main as product_main
    monkeypatch.setattr(product_main, 'load_lm_and_predict', lambda **kw: {'text': ' hello' + str(sorted(kw))})
    if hasattr(product_main.app, 'test_client'):
        r = product_main.app.test_client().get('/inference')
        assert r.status_code == 200 and r.mimetype == 'application/json'
        body = json.loads(r.data)
    else:
        seen = {}
        out = product_main.app({'PATH_INFO': '/inference', 'REQUEST_METHOD': 'GET', 'QUERY_STRING': ''},
                               lambda status, headers: seen.update(status=status, headers=dict(headers)))
        assert seen['status'].startswith('200') and seen['headers']['Content-Type'] == 'application/json'
        assert seen['headers']['Access-Control-Allow-Origin'] == '*'
        body = json.loads(b''.join(out))
        out = product_main.app({'PATH_INFO': '/nope', 'REQUEST_METHOD': 'GET'}, lambda s, h: seen.update(status=s))
        assert seen['status'].startswith('404')
    assert set(body) == {'response'} and set(body['response']) == {'text'}
    assert isinstance(body['response']['text'], str)


def test_local_checkpoint_loading_infers_hparams(tmp_path, monkeypatch):
    import main as product_main
    cfg, sd, d = load_golden('gen_greedy_small')
    root = tmp_path / 'models' / 'rjokes'
    root.mkdir(parents=True)
    torch.save({k: torch.from_numpy(v.copy()) for k, v in sd.items()}, root / 'rjokes.model.pth')
    product_main.pickle_obj([str(s) for s in d['itos']], str(root / 'rjokes.itos.pkl'))
    monkeypatch.setenv('AWD_CACHE_DIR', str(tmp_path))
    monkeypatch.setenv('models_bucket', 'unused-bucket')
    monkeypatch.setattr(product_main, '_MODEL', None)
    m = product_main.load_lm()
    assert (m[0].encoder.num_embeddings, m[0].emb_sz, m[0].n_hid, m[0].n_layers) == (120, 16, 24, 3)
    assert m.itos[0] == 'xxunk' and m.stoi['xxup'] == 4
    assert product_main.load_lm() is m                       # loaded once per process
    monkeypatch.setattr(product_main, '_MODEL', None)
