"""Pins the CPU oracle (oracle/awd_numpy.py, oracle/awd_torch_cpu.py) to the fixtures that
oracle/make_golden.py produced from the live reference module."""
import numpy as np
import pytest
import torch

from oracle import awd_numpy as on
from oracle.awd_torch_cpu import TorchCpuLM
from helpers import FORWARD_FIXTURES, GEN_FIXTURES, load_golden, rel_err


def _replay(cfg, sd, d, dtype):
    p = on.effective_params(sd, cfg, dtype)
    bs, hidden = None, None
    for ci in range(int(d['n_calls'])):
        tok = d[f'call{ci}.tokens']
        if tok.shape[1] != bs:                       # awd_lstm.py:79-81 implicit reset
            bs = tok.shape[1]
            hidden = on.zero_hidden(cfg, bs, dtype)
        dec, raws, hidden = on.forward(p, cfg, tok, hidden)
        yield ci, dec, raws, hidden


@pytest.mark.parametrize('name', FORWARD_FIXTURES)
@pytest.mark.parametrize('dtype', [np.float32, np.float64])
def test_numpy_forward_matches_reference(name, dtype):
    cfg, sd, d = load_golden(name)
    tol = 2e-5 if dtype == np.float32 else 5e-6    # fp64 oracle vs fp32 reference: reference rounding
    for ci, dec, raws, hidden in _replay(cfg, sd, d, dtype):
        assert dec.shape == d[f'call{ci}.decoded'].shape
        assert rel_err(dec, d[f'call{ci}.decoded']) < tol
        for l in range(cfg.n_layers):
            assert rel_err(raws[l], d[f'call{ci}.raw{l}']) < tol
            assert rel_err(hidden[l][0], d[f'call{ci}.h{l}']) < tol
            assert rel_err(hidden[l][1], d[f'call{ci}.c{l}']) < tol


@pytest.mark.parametrize('name', FORWARD_FIXTURES)
def test_torch_port_matches_reference(name):
    cfg, sd, d = load_golden(name)
    m = TorchCpuLM(sd, cfg)
    with torch.no_grad():
        for ci in range(int(d['n_calls'])):
            dec, raws = m.forward(torch.from_numpy(d[f'call{ci}.tokens']))
            assert rel_err(dec.numpy(), d[f'call{ci}.decoded']) < 1e-6
            for l in range(cfg.n_layers):
                assert rel_err(m.hidden[l][1].numpy(), d[f'call{ci}.c{l}']) < 1e-6


def test_state_dict_keys_and_shapes():
    cfg, sd, _ = load_golden('fwd_tiny')
    assert list(sd.keys()) == on.state_dict_keys(cfg)
    assert {k: v.shape for k, v in sd.items()} == on.state_dict_shapes(cfg)
    cfg, sd, _ = load_golden('fwd_l1_nobias')
    assert list(sd.keys()) == on.state_dict_keys(cfg)
    init = on.init_state_dict(cfg, 0)
    assert list(init.keys()) == on.state_dict_keys(cfg)


@pytest.mark.parametrize('name', GEN_FIXTURES)
def test_greedy_generation_matches_reference(name):
    cfg, sd, d = load_golden(name)
    itos = [str(s) for s in d['itos']]
    prompt_words = [str(s) for s in d['prompt_words']]
    p = on.effective_params(sd, cfg)
    stoi = {w: i for i, w in enumerate(itos)}
    prompt = np.array([stoi.get(s, 0) for s in prompt_words], np.int64)[:, None]
    assert prompt[:, 0].tolist() == d['prompt_ids'].tolist()
    ids, gaps = on.generate_ids(p, cfg, prompt, len(d['ids']), on.exclude_tokens(stoi))
    assert ids[:, 0].tolist() == d['ids'].tolist()
    assert gaps.min() > 1e-4
    assert on.sample_model(p, cfg, itos, prompt_words, len(d['ids'])) == str(d['text'])


def test_torch_port_greedy_matches_reference():
    cfg, sd, d = load_golden('gen_greedy_small')
    itos = [str(s) for s in d['itos']]
    stoi = {w: i for i, w in enumerate(itos)}
    m = TorchCpuLM(sd, cfg)
    with torch.no_grad():
        ids = m.generate(torch.from_numpy(d['prompt_ids'])[:, None], len(d['ids']), on.exclude_tokens(stoi), greedy=True)
    assert ids[:, 0].tolist() == d['ids'].tolist()


def test_sampling_law_matches_reference_histogram():
    """first_allowed_law == the empirical law of the unpatched main.py:61-68 (chi-square)."""
    d = np.load(__import__('os').path.join(__import__('helpers').GOLDEN, 'sample_law.npz'))
    itos = [str(s) for s in d['itos']]
    banned = on.exclude_tokens({w: i for i, w in enumerate(itos)})
    law = on.first_allowed_law(d['z'], banned)
    counts = d['counts'].astype(np.float64)
    assert counts[[0] + banned].sum() == 0
    n = counts.sum()
    keep = law > 0
    chi2 = (((counts - n * law) ** 2)[keep] / (n * law[keep])).sum()
    dof = keep.sum() - 1
    assert chi2 < dof + 5 * np.sqrt(2 * dof), (chi2, dof)
    # and the oracle's own exponential-race sampler follows the same law
    cfg = on.LMConfig(len(itos), 4, 4, 1)
    rng = np.random.default_rng(0)
    c2 = np.zeros(len(itos))
    for _ in range(4000):
        keys = d['z'].astype(np.float64) - np.log(rng.exponential(size=len(itos)))
        c2[on.pick_first_allowed(on.topk_order(keys, 10), banned)] += 1
    chi2b = (((c2 - 4000 * law) ** 2)[keep] / (4000 * law[keep])).sum()
    assert chi2b < dof + 5 * np.sqrt(2 * dof)


def test_format_tokens_rules():
    # main.py:73-79: capitalise after '.', '!' and newline; no space before no_space_words
    s = on.format_tokens(['hello'], ['world', '.', 'next', "'s", '!', 'x', '\n', 'y'])
    assert s == " hello world. Next's! X \n Y"
