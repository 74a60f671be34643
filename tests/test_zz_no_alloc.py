"""awd_decode_sample_static -- the entry point for callers that cannot allocate device memory at
decode time (a server whose arena owns the whole GPU; a latency-critical decode thread that
cannot absorb the device-wide stall of cudaMalloc/cudaFree).  It keeps the logits between the
decoder GEMM and the sampler in a fixed 16 MiB array inside the library instead of a workspace.

Checked here through the C ABI, twice in a row on the shared scratch, against the regular pair
awd_decode + awd_sample (ids must be identical: same kernels, same race keys), plus the size
limit and the no-allocation property.  Collected last on purpose (file name), and independent of
what ran before it.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _engine(V=1000, E=400, H=64, L=2, seed=4):
    from awd_b200 import Engine
    rng = np.random.default_rng(seed)
    sd = {}
    enc = rng.uniform(-0.5, 0.5, (V, E)).astype(np.float32)
    sd['0.encoder.weight'] = sd['0.encoder_dp.emb.weight'] = sd['1.decoder.weight'] = enc
    for l in range(L):
        i, h = (E if l == 0 else H), (H if l != L - 1 else E)
        p = f'0.rnns.{l}.'
        sd[p + 'weight_hh_l0_raw'] = sd[p + 'module.weight_hh_l0'] = rng.uniform(-.1, .1, (4 * h, h)).astype(np.float32)
        sd[p + 'module.weight_ih_l0'] = rng.uniform(-.1, .1, (4 * h, i)).astype(np.float32)
        sd[p + 'module.bias_ih_l0'] = np.zeros(4 * h, np.float32)
        sd[p + 'module.bias_hh_l0'] = np.zeros(4 * h, np.float32)
    sd['1.decoder.bias'] = rng.normal(0, 0.1, V).astype(np.float32)
    e = Engine(V, E, H, L)
    e.load_state_dict(sd)
    return e


def _static_call(e, x, sp, step, out):
    from awd_b200 import _lib
    s, keep = sp.to_c()
    return e.lib.awd_decode_sample_static(e._h, C.c_void_p(x.data_ptr()), x.shape[0], C.byref(s), step,
                                          C.c_void_p(out.data_ptr()), e._stream())


def test_static_scratch_entry_matches_regular_path_twice_and_allocates_nothing():
    from awd_b200 import SamplingParams, _lib
    e = _engine()
    g = torch.Generator().manual_seed(0)
    out = torch.empty(64, dtype=torch.int64, device='cuda')
    # two calls in a row on the shared array: a wide batch first, then a narrower one whose
    # rows must not be influenced by what the first call left behind
    cases = [(48, SamplingParams(banned=[3, 4, 5], seed=11), 0), (5, SamplingParams(greedy=True, banned=[3, 4, 5]), 3),
             (17, SamplingParams(temperature=0.7, n_draws=2, banned=list(range(1, 30)), seed=5), 9)]
    free_before = None
    for n, (bs, sp, step) in enumerate(cases):
        x = (torch.randn(bs, 400, generator=g) * 2).cuda()
        want = e.sample(e.decode(x), sp, step)                   # existing entry points
        torch.cuda.synchronize()
        if n == 1:                                               # set-up happened during call 0
            free_before = torch.cuda.mem_get_info()[0]
            torch_before = torch.cuda.memory_allocated()
        _lib.check(_static_call(e, x, sp, step, out))
        torch.cuda.synchronize()
        if n == 1:
            assert torch.cuda.mem_get_info()[0] == free_before, 'device memory was allocated'
            assert torch.cuda.memory_allocated() == torch_before
        assert torch.equal(out[:bs], want), f'case {n}: ids differ from awd_decode + awd_sample'
    e.check_errors()


def test_static_scratch_entry_rejects_what_it_cannot_hold():
    from awd_b200 import SamplingParams, _lib
    e = _engine()
    limit = (16 << 20) // (1000 * 4)                               # streams that fit at V=1000
    out = torch.empty(limit + 1, dtype=torch.int64, device='cuda')
    x = torch.zeros(limit + 1, 400, device='cuda')
    _lib.check(_static_call(e, x[:limit], SamplingParams(greedy=True), 0, out))
    with pytest.raises(ValueError, match='exceeds the fixed'):
        _lib.check(_static_call(e, x, SamplingParams(greedy=True), 0, out))
    with pytest.raises(ValueError, match='16-byte aligned'):      # a row pitch it would have to re-pack
        _lib.check(_static_call(e, x.view(-1)[1:401].view(1, 400), SamplingParams(greedy=True), 0, out))
    torch.cuda.synchronize()
