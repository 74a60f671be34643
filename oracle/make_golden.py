"""Generate tests/golden/*.npz from the LIVE reference (run in the build container only).

    python oracle/make_golden.py

Imports ``/root/reference/pytorch_models/awd_lstm.py`` unmodified and exec's
``/root/reference/main.py`` with (a) the two stray ':' at main.py:90 and :98 removed so that it
parses and (b) ``flask``/``flask_cors``/``boto3``/``requests`` stubbed (not installed here).  The
functions used (``sample_model`` main.py:40-81) are otherwise untouched; the greedy fixtures
replace the one expression ``torch.multinomial(res[-1].exp(), 10)`` (main.py:61) with
``torch.topk(res[-1], 10).indices`` as SURVEY.md §8c prescribes.  The reference cannot travel to
the GPU box, so the fixtures are committed; this script is the provenance.
"""
import os
import sys
import types

import numpy as np
import torch

REF = '/root/reference'
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'tests', 'golden')
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))


def load_reference():
    sys.path.insert(0, REF)
    from pytorch_models import awd_lstm as ref_awd  # noqa
    sys.path.pop(0)
    for name in ('flask', 'flask_cors', 'boto3', 'requests'):
        if name not in sys.modules:
            m = types.ModuleType(name)
            sys.modules[name] = m
    fl = sys.modules['flask']

    class _Flask:
        def __init__(self, *a, **k): pass
        def route(self, *a, **k): return lambda f: f
    fl.Flask = _Flask
    fl.Response = lambda **k: k
    sys.modules['flask_cors'].CORS = lambda app: None
    sys.modules['flask_cors'].cross_origin = lambda *a, **k: (lambda f: f)
    src = open(os.path.join(REF, 'main.py')).read().split('\n')
    assert src[89].rstrip().endswith('):') and src[97].rstrip().endswith('):')
    src[89] = src[89].rstrip()[:-1]
    src[97] = src[97].rstrip()[:-1]
    src = '\n'.join(src)
    ns_sample = {'__name__': 'ref_main'}
    sys.modules['pytorch_models'] = sys.modules['pytorch_models']
    exec(compile(src, 'main.py', 'exec'), ns_sample)
    greedy_src = src.replace('torch.multinomial(res[-1].exp(), 10)', 'torch.topk(res[-1], 10).indices')
    assert greedy_src != src
    ns_greedy = {'__name__': 'ref_main_greedy'}
    exec(compile(greedy_src, 'main.py', 'exec'), ns_greedy)
    return ref_awd, ns_sample, ns_greedy


def sd_to_np(sd):
    return {k: v.detach().cpu().numpy().copy() for k, v in sd.items()}


def build(ref_awd, V, E, H, L, seed, tie=True, bias=True, bias_noise=0.0):
    torch.manual_seed(seed)
    m = ref_awd.get_language_model(vocab_sz=V, emb_sz=E, n_hid=H, n_layers=L, pad_token=1, tie_weights=tie, bias=bias)
    if bias and bias_noise:
        with torch.no_grad():
            m[1].decoder.bias.normal_(0, bias_noise)
    m.eval()
    m.reset()
    return m


def run_calls(m, calls):
    """calls: list of token arrays; returns per-call outputs (state carried as the module does)."""
    out = {}
    with torch.no_grad():
        for ci, tok in enumerate(calls):
            dec, raws, outs = m(torch.from_numpy(tok))
            out[f'call{ci}.tokens'] = tok
            out[f'call{ci}.decoded'] = dec.numpy().copy()
            for l, r in enumerate(raws):
                out[f'call{ci}.raw{l}'] = r.numpy().copy()
                assert torch.equal(r, outs[l])
            for l, (h, c) in enumerate(m[0].hidden):
                out[f'call{ci}.h{l}'] = h.numpy().copy()
                out[f'call{ci}.c{l}'] = c.numpy().copy()
    return out


def fixture_forward(ref_awd, name, V, E, H, L, seed, shapes, tie=True, bias=True, bias_noise=0.05):
    m = build(ref_awd, V, E, H, L, seed, tie, bias, bias_noise)
    g = torch.Generator().manual_seed(seed + 1)
    calls = [torch.randint(0, V, s, generator=g).numpy() for s in shapes]
    data = {'cfg': np.array([V, E, H, L, int(tie), int(bias)], np.int64)}
    for k, v in sd_to_np(m.state_dict()).items():
        data['sd.' + k] = v
    data.update(run_calls(m, calls))
    data['n_calls'] = np.array(len(calls))
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **data)
    print('wrote', name, {k: v.shape for k, v in data.items() if k.startswith('call0')})


def fixture_forward_seeded(ref_awd, name, V, E, H, L, seed, shapes):
    """Shipped layer sizes: the weights (80 MB) are NOT stored; they come from
    oracle.awd_numpy.init_state_dict(cfg, seed) (numpy PCG64, stable across machines), are loaded
    into the live reference here, and only the outputs are committed."""
    from oracle.awd_numpy import LMConfig, init_state_dict
    cfg = LMConfig(V, E, H, L)
    sd = init_state_dict(cfg, seed, decoder_bias_scale=0.05)
    m = build(ref_awd, V, E, H, L, 0)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in sd.items()})
    m.eval(); m.reset()
    g = torch.Generator().manual_seed(seed + 1)
    calls = [torch.randint(0, V, s, generator=g).numpy() for s in shapes]
    data = {'cfg': np.array([V, E, H, L, 1, 1], np.int64), 'init_seed': np.array(seed),
            'decoder_bias_scale': np.array(0.05)}
    data.update(run_calls(m, calls))
    data['n_calls'] = np.array(len(calls))
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **data)
    print('wrote', name)


def fixture_alias(ref_awd, name):
    """A checkpoint whose aliased keys hold DIFFERENT values: pins the last-key-wins rule of
    load_state_dict (main.py:99) that SURVEY.md §5.4 describes."""
    V, E, H, L = 30, 8, 12, 3
    m = build(ref_awd, V, E, H, L, 7)
    sd = sd_to_np(m.state_dict())
    rng = np.random.default_rng(3)
    for k in list(sd):
        sd[k] = (sd[k] + rng.normal(0, 0.05, sd[k].shape)).astype(np.float32)   # every key independent now
    m2 = build(ref_awd, V, E, H, L, 8)
    m2.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    m2.eval(); m2.reset()
    tok = np.random.default_rng(4).integers(0, V, (4, 2))
    data = {'cfg': np.array([V, E, H, L, 1, 1], np.int64)}
    for k, v in sd.items():
        data['sd.' + k] = v
    data.update(run_calls(m2, [tok]))
    data['n_calls'] = np.array(1)
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **data)
    print('wrote', name)


def fixture_generate(ref_awd, ns_greedy, name, V, E, H, L, seed, prompt_words, l):
    from oracle.awd_numpy import synthetic_itos
    m = build(ref_awd, V, E, H, L, seed, bias_noise=0.0)
    # scale the weights up: the recurrence turns chaotic (varied tokens instead of one fixed
    # point) and the top-1/top-2 logit gaps stay far above fp32 reduction noise
    with torch.no_grad():
        m[0].encoder.weight.mul_(20.0)
        for r in m[0].rnns:
            r.weight_hh_l0_raw.mul_(8.0)
            r.module.weight_ih_l0.mul_(8.0)
    m.reset()
    itos = synthetic_itos(V)
    m.itos = itos
    m.stoi = {w: i for i, w in enumerate(itos)}
    fed = []
    orig_forward = m.forward

    def rec_forward(x):
        fed.append(int(x[0, 0]))
        return orig_forward(x)
    m.forward = rec_forward
    with torch.no_grad():
        text = ns_greedy['sample_model'](m, prompt_words, l=l)
    ids = np.array(fed[len(prompt_words):], np.int64)
    assert len(ids) == l
    data = {'cfg': np.array([V, E, H, L, 1, 1], np.int64), 'ids': ids,
            'prompt_ids': np.array(fed[:len(prompt_words)], np.int64),
            'text': np.array(text), 'prompt_words': np.array(prompt_words), 'itos': np.array(itos)}
    for k, v in sd_to_np(m.state_dict()).items():
        data['sd.' + k] = v
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **data)
    print('wrote', name, repr(text[:80]))


def fixture_sample_law(ns_sample, name, V=24, n=40000):
    """Empirical law of the UNPATCHED sampling rule (main.py:61-68, real torch.multinomial) on a
    fixed logit vector: a stub model returns the same logits every step, sample_model runs n
    steps, and the histogram of the ids it feeds back is committed."""
    from oracle.awd_numpy import synthetic_itos
    itos = synthetic_itos(V)
    z = torch.from_numpy(np.random.default_rng(5).normal(0, 1.5, V).astype(np.float32))
    fed = []

    class _Core:
        bs = 1

    class _Stub:
        def __init__(self):
            self.itos, self.stoi, self.core = itos, {w: i for i, w in enumerate(itos)}, _Core()
        def __getitem__(self, i): return self.core
        def eval(self): pass
        def reset(self): pass
        def forward(self, x):
            fed.append(int(x[0, 0]))
            return (z[None, :],)
    torch.manual_seed(123)
    ns_sample['sample_model'](_Stub(), [''], l=n)
    counts = np.bincount(np.array(fed[1:]), minlength=V)
    np.savez_compressed(os.path.join(OUT, name + '.npz'), z=z.numpy(), counts=counts, itos=np.array(itos))
    print('wrote', name, counts[:8])


def check_torch_port(ref_awd):
    """The torch-CPU port used for the CPU baseline must equal the live reference."""
    from oracle.awd_numpy import LMConfig
    from oracle.awd_torch_cpu import TorchCpuLM
    m = build(ref_awd, 200, 16, 24, 3, 11)
    port = TorchCpuLM(sd_to_np(m.state_dict()), LMConfig(200, 16, 24, 3))
    tok = torch.randint(0, 200, (6, 5), generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        a = m(tok)[0]
        b = port.forward(tok)[0]
        a2 = m(tok)[0]
        b2 = port.forward(tok)[0]
    assert torch.allclose(a, b, atol=1e-6) and torch.allclose(a2, b2, atol=1e-6), 'torch port != reference'
    print('torch-CPU port matches the live reference (max abs', float((a2 - b2).abs().max()), ')')


def main():
    os.makedirs(OUT, exist_ok=True)
    ref_awd, ns_sample, ns_greedy = load_reference()
    check_torch_port(ref_awd)
    # state carry (call1 continues call0), implicit reset on batch change (call2), single step (call3)
    fixture_forward(ref_awd, 'fwd_tiny', 40, 8, 12, 3, 1, [(5, 3), (3, 3), (2, 2), (1, 2)])
    # odd sizes: nothing a multiple of 4/16/64; untied; two layers
    fixture_forward(ref_awd, 'fwd_odd', 301, 20, 46, 3, 2, [(7, 4), (1, 4)])
    fixture_forward(ref_awd, 'fwd_untied_l2', 97, 12, 18, 2, 3, [(4, 2)], tie=False)
    fixture_forward(ref_awd, 'fwd_l1_nobias', 64, 16, 16, 1, 4, [(3, 5)], bias=False)
    # shipped layer shapes at a reduced vocabulary (keeps the file small): E=400,H=1150
    fixture_forward_seeded(ref_awd, 'fwd_shipped_dims', 500, 400, 1150, 3, 5, [(3, 2), (1, 2)])
    fixture_alias(ref_awd, 'fwd_alias')
    fixture_sample_law(ns_sample, 'sample_law')
    fixture_generate(ref_awd, ns_greedy, 'gen_greedy_small', 120, 16, 24, 3, 6, ['w20', 'w31', 'zzz-oov', '.'], 40)
    fixture_generate(ref_awd, ns_greedy, 'gen_greedy_empty_prompt', 80, 12, 20, 3, 7, [''], 25)


if __name__ == '__main__':
    main()
