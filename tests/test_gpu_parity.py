"""Parity of the CUDA path (through the C ABI, via awd_b200.Engine / the drop-in module) with
the CPU oracle and the golden fixtures produced from the live reference.

Tolerances (BASELINE.json north_star): fp32 mode -- logits and hidden/cell states within 1e-4
relative (max|a-b| / max|b|), identical greedy token sequences, identical gather rows.
"""
import numpy as np
import pytest
import torch

from helpers import FORWARD_FIXTURES, GEN_FIXTURES, load_golden, rel_err
from oracle import awd_numpy as on

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-4


def make_engine(cfg, sd, precision='fp32'):
    from awd_b200 import Engine
    e = Engine(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, tie_weights=cfg.tie_weights, bias=cfg.bias,
               precision=precision)
    e.load_state_dict(sd)
    return e


def test_native_library_is_the_one_running():
    from awd_b200 import _lib
    lib = _lib.load()
    assert b'sm_100a' in lib.awd_version()
    with open('/proc/self/maps') as f:
        assert 'libawd_b200.so' in f.read()


@pytest.mark.parametrize('name', FORWARD_FIXTURES)
def test_engine_forward_matches_reference_fixture(name):
    cfg, sd, d = load_golden(name)
    e = make_engine(cfg, sd)
    n0 = e.launch_count()
    for ci in range(int(d['n_calls'])):
        tok = torch.from_numpy(d[f'call{ci}.tokens'])
        logits, raws = e.forward(tok)
        assert rel_err(logits.cpu().numpy(), d[f'call{ci}.decoded']) < TOL_FP32
        hidden = e.get_hidden()
        for l in range(cfg.n_layers):
            assert raws[l].shape == d[f'call{ci}.raw{l}'].shape
            assert rel_err(raws[l].cpu().numpy(), d[f'call{ci}.raw{l}']) < TOL_FP32
            assert rel_err(hidden[l][0].cpu().numpy(), d[f'call{ci}.h{l}']) < TOL_FP32
            assert rel_err(hidden[l][1].cpu().numpy(), d[f'call{ci}.c{l}']) < TOL_FP32
    assert e.launch_count() > n0


@pytest.mark.parametrize('name', ['fwd_tiny', 'fwd_alias', 'fwd_untied_l2', 'fwd_l1_nobias'])
def test_dropin_module_matches_reference_fixture(name):
    """Same calls a user of the reference makes: get_language_model, strict load_state_dict,
    eval, reset, model(x); parameters and inputs on the CPU as in main.py."""
    from pytorch_models import awd_lstm
    cfg, sd, d = load_golden(name)
    m = awd_lstm.get_language_model(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, pad_token=1,
                                    tie_weights=cfg.tie_weights, bias=cfg.bias)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in sd.items()})
    m.eval()
    m.reset()
    for ci in range(int(d['n_calls'])):
        dec, raws, outs = m(torch.from_numpy(d[f'call{ci}.tokens']))
        assert dec.device.type == 'cpu' and tuple(dec.shape) == d[f'call{ci}.decoded'].shape
        assert rel_err(dec.numpy(), d[f'call{ci}.decoded']) < TOL_FP32
        assert m[0].bs == d[f'call{ci}.tokens'].shape[1]
        for l in range(cfg.n_layers):
            assert rel_err(raws[l].numpy(), d[f'call{ci}.raw{l}']) < TOL_FP32
            assert outs[l] is raws[l]
            h, c = m[0].hidden[l]
            assert rel_err(h.numpy(), d[f'call{ci}.h{l}']) < TOL_FP32
            assert rel_err(c.numpy(), d[f'call{ci}.c{l}']) < TOL_FP32


def test_dropin_module_split_halves_and_reset():
    """model[0](x) then model[1](...) equals model(x); reset() zeroes the state; editing a
    parameter in place is picked up."""
    from pytorch_models import awd_lstm
    cfg, sd, d = load_golden('fwd_tiny')
    m = awd_lstm.get_language_model(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, pad_token=1)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in sd.items()})
    m.eval(); m.reset()
    tok = torch.from_numpy(d['call0.tokens'])
    dec, _, _ = m(tok)
    m.reset()
    assert float(m[0].hidden[0][0].abs().max()) == 0.0
    raws, outs = m[0](tok)
    dec2, _, _ = m[1]((raws, outs))
    assert torch.allclose(dec, dec2, atol=1e-6)
    emb = m[0].encoder_dp(tok)
    assert np.array_equal(emb.numpy(), sd['0.encoder.weight'][d['call0.tokens']])
    with torch.no_grad():
        m[1].decoder.bias.add_(1.0)
    m.reset()
    dec3, _, _ = m(tok)
    assert torch.allclose(dec3, dec + 1.0, atol=1e-5)
    m.train()
    with pytest.raises(NotImplementedError):
        m(tok)


def test_gather_is_exact_and_checks_range():
    cfg = on.LMConfig(1000, 400, 1150, 3)
    sd = on.init_state_dict(cfg, 3)
    e = make_engine(cfg, sd)
    tok = torch.randint(0, 1000, (7, 33), generator=torch.Generator().manual_seed(0))
    out = e.gather(tok).cpu().numpy()
    assert np.array_equal(out, sd['0.encoder.weight'][tok.numpy()])
    out = e.gather(tok.cuda()).cpu().numpy()
    assert np.array_equal(out, sd['0.encoder.weight'][tok.numpy()])
    with pytest.raises(IndexError):
        e.forward(torch.tensor([[1000]]))
    with pytest.raises(IndexError):
        e.forward(torch.tensor([[-1]]))
    e.forward(torch.tensor([[999]]))


@pytest.mark.parametrize('bs,sl', [(1, 4), (3, 2), (8, 3), (9, 2), (70, 2), (300, 3)])
def test_engine_forward_vs_oracle_shipped_dims(bs, sl):
    """Shipped layer sizes (E=400,H=1150,L=3) at batch sizes that exercise every kernel
    variant: skinny (1,2,4,8 streams), 64x64 and 128x128 tiles, single- and multi-step calls."""
    cfg = on.LMConfig(1200, 400, 1150, 3)
    sd = on.init_state_dict(cfg, 9, decoder_bias_scale=0.05)
    e = make_engine(cfg, sd)
    p = on.effective_params(sd, cfg)
    g = torch.Generator().manual_seed(bs * 100 + sl)
    hidden = on.zero_hidden(cfg, bs)
    for call in range(2):      # second call continues from the carried state; a 1-step call last
        T = sl if call == 0 else 1
        tok = torch.randint(0, cfg.vocab_sz, (T, bs), generator=g)
        ref_dec, ref_raws, hidden = on.forward(p, cfg, tok.numpy(), hidden)
        logits, raws = e.forward(tok.cuda())
        assert rel_err(logits.cpu().numpy(), ref_dec) < TOL_FP32
        for l in range(cfg.n_layers):
            assert rel_err(raws[l].cpu().numpy(), ref_raws[l]) < TOL_FP32
        hh = e.get_hidden()
        for l in range(cfg.n_layers):
            assert rel_err(hh[l][1].cpu().numpy(), hidden[l][1]) < TOL_FP32
    last, _ = e.forward(tok.cuda(), last_step_only=True)
    assert last.shape == (bs, cfg.vocab_sz)


def test_state_set_get_roundtrip_and_explicit_hidden():
    cfg, sd, d = load_golden('fwd_odd')
    e = make_engine(cfg, sd)
    e.reset(4)
    hs = [(torch.randn(1, 4, H), torch.randn(1, 4, H)) for H in e.layer_hidden]
    e.set_hidden(hs)
    back = e.get_hidden()
    for (h, c), (h2, c2) in zip(hs, back):
        assert torch.equal(h, h2.cpu()) and torch.equal(c, c2.cpu())
    # forward from an explicit state == oracle from the same state
    p = on.effective_params(sd, cfg)
    tok = d['call0.tokens'][:2]
    ref_dec, _, _ = on.forward(p, cfg, tok, [(h.numpy(), c.numpy()) for h, c in hs])
    logits, _ = e.forward(torch.from_numpy(tok))
    assert rel_err(logits.cpu().numpy(), ref_dec) < TOL_FP32


@pytest.mark.parametrize('name', GEN_FIXTURES)
def test_greedy_generation_identical_to_reference(name):
    """fp32 greedy decoding reproduces the reference's patched sample_model token for token, and
    the product's sample_model reproduces its text."""
    import main as product_main
    from pytorch_models import awd_lstm
    cfg, sd, d = load_golden(name)
    itos = [str(s) for s in d['itos']]
    m = awd_lstm.get_language_model(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, pad_token=1)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in sd.items()})
    m.itos, m.stoi = itos, {w: i for i, w in enumerate(itos)}
    text = product_main.sample_model(m, [str(s) for s in d['prompt_words']], l=len(d['ids']), greedy=True)
    assert text == str(d['text'])
    from awd_b200 import SamplingParams
    from awd_b200.text import exclude_token_ids
    sp = SamplingParams(greedy=True, banned=exclude_token_ids(m.stoi))
    ids = m.generate_ids(torch.from_numpy(d['prompt_ids'])[:, None], len(d['ids']), sp)
    assert ids[:, 0].tolist() == d['ids'].tolist()


def test_greedy_generation_batch_vs_oracle_and_batch_sharding():
    """8 streams x 24 tokens at shipped layer sizes: identical ids to the fp32 oracle (min top-1/
    top-2 gap reported), and shard k of n reproduces columns [k*B/n,(k+1)*B/n) of the full run."""
    from awd_b200 import SamplingParams
    cfg = on.LMConfig(600, 400, 1150, 3)
    sd = on.init_state_dict(cfg, 21)
    for k in sd:
        if 'weight_ih' in k or 'weight_hh' in k:
            sd[k] = sd[k] * 6.0          # chaotic regime: varied tokens, wide argmax gaps
    sd['0.encoder.weight'] = sd['0.encoder_dp.emb.weight'] = sd['1.decoder.weight'] = sd['1.decoder.weight'] * 20.0
    e = make_engine(cfg, sd)
    p = on.effective_params(sd, cfg)
    prompt = torch.randint(6, 600, (2, 8), generator=torch.Generator().manual_seed(5))
    banned = [3, 4, 5]
    ref_ids, gaps = on.generate_ids(p, cfg, prompt.numpy(), 24, banned)
    sp = SamplingParams(greedy=True, banned=banned)
    ids = e.generate(prompt, 24, sp).cpu().numpy()
    mism = np.argwhere(ids != ref_ids)
    if len(mism):      # SURVEY.md 7 hard part 3: a mismatch is only acceptable at a near-tie
        t, b = mism[0]
        assert gaps[t, b] < 1e-5, f'diverged at step {t} stream {b} with gap {gaps[t, b]}'
    else:
        assert len(set(ids.flatten().tolist())) > 10
    for k in range(4):
        part = e.generate(prompt[:, 2 * k:2 * k + 2].contiguous(), 24, sp).cpu().numpy()
        assert np.array_equal(part, ids[:, 2 * k:2 * k + 2])


# ------------------------------------------------------------------------------------ sampling
def _philox4x32_10(c, k):
    c = [np.uint64(x) for x in c]
    k0, k1 = np.uint64(k[0]), np.uint64(k[1])
    M0, M1, mask = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0xffffffff)
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        c = [((p1 >> np.uint64(32)) ^ c[1] ^ k0) & mask, p1 & mask, ((p0 >> np.uint64(32)) ^ c[3] ^ k1) & mask, p0 & mask]
        k0 = (k0 + np.uint64(0x9E3779B9)) & mask
        k1 = (k1 + np.uint64(0xBB67AE85)) & mask
    return [int(x) for x in c]


def _host_race_pick(z, b, step, seed, banned_with_zero, n_draws, temperature=1.0):
    V = len(z)
    keys = np.empty(V)
    for c4 in range((V + 3) // 4):
        r = _philox4x32_10([c4, b, step & 0xffffffff, step >> 32], [seed & 0xffffffff, seed >> 32])
        for q in range(4):
            i = c4 * 4 + q
            if i < V:
                u = ((r[q] >> 9) + 0.5) / 8388608.0
                keys[i] = float(z[i]) / temperature - np.log(-np.log(u))
    order = np.lexsort((np.arange(V), -keys))[:n_draws]
    for r in order:
        if int(r) not in banned_with_zero:
            return int(r), keys
    return int(order[0]), keys


def test_sampler_matches_host_replay_of_the_same_race():
    """The device sampler is an exponential race keyed by Philox4x32-10(column/4, stream, step);
    replaying the same race on the host with the reference's 10-draw / first-allowed rule
    (main.py:61-68) must give the same ids, including the fallback when >= n_draws banned ids lead."""
    from awd_b200 import SamplingParams
    cfg = on.LMConfig(203, 8, 8, 1)
    e = make_engine(cfg, on.init_state_dict(cfg, 0))
    rng = np.random.default_rng(1)
    z = rng.normal(0, 2.0, (6, 203)).astype(np.float32)
    banned = [3, 4, 5]
    for n_draws, bset in [(10, banned), (2, list(range(1, 40)))]:
        z2 = z.copy()
        if n_draws == 2:
            z2[:, :40] += 6.0       # banned ids dominate -> the fallback branch fires for most rows
        sp = SamplingParams(n_draws=n_draws, banned=bset, seed=0x1234567890ab)
        got = e.sample(torch.from_numpy(z2), sp, step=7).cpu().numpy()
        fallbacks = 0
        for b in range(6):
            want, keys = _host_race_pick(z2[b], b, 7, 0x1234567890ab, set([0] + bset), n_draws)
            fallbacks += want in set([0] + bset)
            if got[b] != want:
                s = np.sort(keys)
                assert min(abs(keys[got[b]] - keys[want]), s[-1] - s[-2]) < 1e-4
        if n_draws == 2:
            assert fallbacks >= 3
    g = e.sample(torch.from_numpy(z), SamplingParams(greedy=True, banned=banned), step=0).cpu().numpy()
    for b in range(6):
        assert g[b] == on.pick_first_allowed(on.topk_order(z[b], 10), banned)


def test_sampler_distribution_matches_reference_law():
    """chi-square of 40k device draws against the law pinned to the reference by
    tests/golden/sample_law.npz (the unpatched main.py:61-68 with torch.multinomial)."""
    from awd_b200 import SamplingParams
    import os
    from helpers import GOLDEN
    d = np.load(os.path.join(GOLDEN, 'sample_law.npz'))
    itos = [str(s) for s in d['itos']]
    banned = on.exclude_tokens({w: i for i, w in enumerate(itos)})
    V = len(itos)
    cfg = on.LMConfig(V, 8, 8, 1)
    e = make_engine(cfg, on.init_state_dict(cfg, 0))
    z = torch.from_numpy(np.tile(d['z'], (4000, 1)))
    counts = np.zeros(V)
    for step in range(10):
        ids = e.sample(z, SamplingParams(banned=banned, seed=99), step=step).cpu().numpy()
        counts += np.bincount(ids, minlength=V)
    law = on.first_allowed_law(d['z'], banned)
    n = counts.sum()
    keep = law > 0
    assert counts[~keep].sum() == 0
    chi2 = (((counts - n * law) ** 2)[keep] / (n * law[keep])).sum()
    dof = keep.sum() - 1
    assert chi2 < dof + 5 * np.sqrt(2 * dof), (chi2, dof)
    # and against the reference's own histogram (two-sample chi-square)
    ref = d['counts'].astype(np.float64)
    k1, k2 = np.sqrt(ref.sum() / n), np.sqrt(n / ref.sum())
    chi2b = (((k1 * counts - k2 * ref) ** 2)[keep] / (counts + ref)[keep]).sum()
    assert chi2b < dof + 5 * np.sqrt(2 * dof), (chi2b, dof)


# ------------------------------------------------------------------- full-size property checks
def test_full_size_forward_properties():
    """BASELINE config 2 (bptt 70 x batch 256, V=60000) is too large for the CPU oracle to finish
    in seconds; check size-independent properties instead: (a) splitting the sequence into two
    calls gives the same final state and last-step logits as one call (state carry), (b) any
    single stream run alone reproduces its column of the batch, (c) last_step_only rows equal
    the tail rows of the full logits, (d) a sampled sub-block of logits equals the oracle's
    decoder applied to the engine's own outputs."""
    cfg = on.LMConfig(60000, 400, 1150, 3)
    sd = on.init_state_dict(cfg, 0)
    e = make_engine(cfg, sd)
    tok = torch.randint(0, 60000, (70, 256), generator=torch.Generator().manual_seed(1)).cuda()
    logits, raws = e.forward(tok)
    assert logits.shape == (70 * 256, 60000)
    assert bool(torch.isfinite(logits).all())
    h_full = e.get_hidden()
    tail = logits[-256:].clone()
    sub_rows = torch.tensor([0, 255, 256 * 35 + 7, 70 * 256 - 1])
    sub = logits[sub_rows.cuda()].cpu().numpy()
    out_last = raws[-1].reshape(70 * 256, 400)[sub_rows.cuda()].cpu().numpy()
    del logits
    want = out_last @ sd['1.decoder.weight'].T + sd['1.decoder.bias']
    assert rel_err(sub, want) < TOL_FP32
    e.reset(256)
    e.forward(tok[:35], need_logits=False, need_raw=False)
    last, _ = e.forward(tok[35:], need_raw=False, last_step_only=True)
    assert rel_err(last.cpu().numpy(), tail.cpu().numpy()) < 1e-5
    for (h, c), (h2, c2) in zip(h_full, e.get_hidden()):
        assert rel_err(h2.cpu().numpy(), h.cpu().numpy()) < 1e-5
        assert rel_err(c2.cpu().numpy(), c.cpu().numpy()) < 1e-5
    one, raws1 = e.forward(tok[:, 17:18].contiguous(), last_step_only=True)
    assert rel_err(one.cpu().numpy(), tail[17:18].cpu().numpy()) < 1e-5
    assert rel_err(raws1[0].cpu().numpy(), raws[0][:, 17:18].cpu().numpy()) < 1e-5


def test_errors_follow_the_reference_conventions():
    from awd_b200 import Engine
    from awd_b200._lib import AwdError
    cfg = on.LMConfig(50, 8, 12, 3)
    sd = on.init_state_dict(cfg, 0)
    e = Engine(50, 8, 12, 3)
    with pytest.raises(AwdError):                       # forward before weights
        e.forward(torch.zeros(1, 1, dtype=torch.long))
    bad = dict(sd)
    del bad['0.rnns.1.module.weight_hh_l0']
    with pytest.raises(AwdError, match='missing key'):   # strict load (main.py:99)
        e.load_state_dict(bad)
    with pytest.raises(ValueError, match='size mismatch'):
        e.load_state_dict({'1.decoder.bias': np.zeros(51, np.float32)})
    with pytest.raises(ValueError, match='unexpected key'):
        e.load_state_dict({'2.nope': np.zeros(1, np.float32)})
    e.load_state_dict(sd)
    with pytest.raises(ValueError):
        e.forward(torch.zeros(3, dtype=torch.long))


@pytest.mark.parametrize('bs,V', [(40, 1003), (300, 5000)])
def test_fp32_fused_decoder_race_equals_decoder_then_sampler(bs, V, monkeypatch):
    """fp32 generation with more than 8 streams can fuse the sampling race into the decoder GEMM's epilogue (AWD_F32_FUSE=1: the
    logits never reach HBM; off by default because the two-kernel form measured faster).  The keys are those of the stand-alone sampler -- Philox(column / 4, stream, step) -- so the ids must EQUAL what
    awd_forward's logits + awd_sample give step by step with the same seed: sampled, greedy, and with so many disallowed ids that
    the "else the first candidate" rule fires (reference main.py:61-68)."""
    from awd_b200 import SamplingParams
    cfg = on.LMConfig(V, 24, 40, 2)
    sd = on.init_state_dict(cfg, 17, decoder_bias_scale=0.5)
    sd['1.decoder.bias'][:30] += 4.0
    e = make_engine(cfg, sd)
    monkeypatch.setenv('AWD_F32_FUSE', '1')
    prompt = torch.randint(6, V, (3, bs), generator=torch.Generator().manual_seed(bs))
    n_new = 5
    for sp in (SamplingParams(banned=[3, 4, 5], seed=0xabcdef123), SamplingParams(greedy=True, banned=[3, 4, 5]),
               SamplingParams(n_draws=2, banned=list(range(1, 25)), seed=99, temperature=0.7)):
        launches = e.launch_count()
        got = e.generate(prompt, n_new, sp).cpu().numpy()
        fused_launches = e.launch_count() - launches
        e.reset(bs)
        z = e.forward(prompt, last_step_only=True, need_raw=False)[0]
        want = []
        for i in range(n_new):
            ids = e.sample(z, sp, step=i)
            want.append(ids.cpu().numpy())
            z = e.forward(ids[None, :], last_step_only=True, need_raw=False)[0]
        want = np.stack(want)
        assert np.array_equal(got, want), (bs, sp, int((got != want).sum()))
        if sp.n_draws == 2:
            assert np.isin(got, [0] + list(range(1, 25))).mean() > 0.2          # the fallback branch really fired
