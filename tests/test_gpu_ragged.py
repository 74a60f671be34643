"""Ragged prompts (awd_generate_ragged): requests with different prompt lengths share one device batch; every
stream must produce what it produces alone (reference main.py:54-57 feeds each request only its own words)."""
import numpy as np
import pytest
import torch

from oracle import awd_numpy as on

pytestmark = pytest.mark.gpu


def _model(seed=41):
    cfg = on.LMConfig(700, 400, 1150, 3)
    sd = on.init_state_dict(cfg, seed, decoder_bias_scale=0.05)
    for k in sd:
        if 'weight_ih' in k or 'weight_hh' in k:
            sd[k] = sd[k] * 4.0
    sd['0.encoder.weight'] = sd['0.encoder_dp.emb.weight'] = sd['1.decoder.weight'] = sd['1.decoder.weight'] * 20.0
    return cfg, sd


@pytest.mark.parametrize('precision,bs', [('fp32', 5), ('fp32', 12), ('bf16', 6), ('bf16', 20)])
def test_ragged_batch_equals_streams_run_alone(precision, bs):
    from awd_b200 import Engine, SamplingParams
    cfg, sd = _model()
    e = Engine(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, precision=precision)
    e.load_state_dict(sd)
    rng = np.random.default_rng(bs)
    P = 5
    lens = [int(x) for x in rng.integers(1, P + 1, bs)]
    lens[0], lens[1] = P, 1
    full = rng.integers(6, 700, (P, bs))
    prompt = np.zeros((P, bs), np.int64)
    for b in range(bs):
        prompt[P - lens[b]:, b] = full[P - lens[b]:, b]          # right-aligned; rows above are ignored
        prompt[:P - lens[b], b] = 699                             # garbage that must not matter
    sp = SamplingParams(greedy=True, banned=[3, 4, 5])
    got = e.generate(torch.from_numpy(prompt), 8, sp, prompt_len=lens).cpu().numpy()
    p = on.effective_params(sd, cfg)
    agree = []
    for b in range(bs):
        if precision == 'fp32':                                   # exact: the oracle on this stream alone
            want, _ = on.generate_ids(p, cfg, full[P - lens[b]:, b:b + 1], 8, [3, 4, 5])
            want = want[:, 0]
        else:       # bf16: the engine on this stream alone, through the SAME kernels (few-stream kernel for bs <= 8,
                    # tcgen05 path otherwise: there the lone stream is replicated 9 times to stay on that path)
            reps = 1 if bs <= 8 else 9
            alone = np.ascontiguousarray(np.repeat(full[P - lens[b]:, b:b + 1], reps, axis=1))
            want = e.generate(torch.from_numpy(alone), 8, sp).cpu().numpy()[:, 0]
        agree.append(int(np.argmax(np.append(got[:, b] != want, True))))
    if precision == 'fp32':
        assert min(agree) == 8, agree
    else:                                                         # same kernels, same per-row arithmetic
        assert min(agree) == 8, agree


def test_ragged_rejects_bad_lengths_and_uniform_equals_plain():
    from awd_b200 import Engine, SamplingParams
    cfg, sd = _model(43)
    e = Engine(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, precision='bf16')
    e.load_state_dict(sd)
    prompt = torch.randint(6, 700, (3, 10), generator=torch.Generator().manual_seed(0))
    sp = SamplingParams(greedy=True, banned=[3, 4, 5])
    with pytest.raises(ValueError):
        e.generate(prompt, 2, sp, prompt_len=[0] * 10)
    with pytest.raises(ValueError):
        e.generate(prompt, 2, sp, prompt_len=[4] * 10)
    a = e.generate(prompt, 4, sp, prompt_len=[3] * 10).cpu().numpy()
    b = e.generate(prompt, 4, sp).cpu().numpy()
    assert np.array_equal(a, b)
