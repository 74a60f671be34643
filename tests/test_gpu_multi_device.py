"""Two replicas in ONE process (the serving topology of awd_b200.scheduler.BatchScheduler): skipped with
fewer than two GPUs.  Each replica must give what a single-GPU run gives for the same streams."""
import numpy as np
import pytest
import torch

from oracle import awd_numpy as on

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
@pytest.mark.parametrize('precision', ['fp32', 'bf16'])
def test_two_engines_two_devices_one_process(precision):
    import threading
    from awd_b200 import Engine, SamplingParams
    cfg = on.LMConfig(900, 400, 1150, 3)
    sd = on.init_state_dict(cfg, 4, decoder_bias_scale=0.05)
    engines = [Engine(900, 400, 1150, 3, precision=precision, device=d) for d in range(2)]
    for e in engines:
        e.load_state_dict(sd)
    prompt = torch.randint(6, 900, (2, 300), generator=torch.Generator().manual_seed(2))
    sp = SamplingParams(greedy=True, banned=[3, 4, 5])
    out = [None, None]

    def run(d):
        torch.cuda.set_device(d)
        out[d] = engines[d].generate(prompt[:, 150 * d:150 * (d + 1)].contiguous(), 6, sp).cpu().numpy()
    ts = [threading.Thread(target=run, args=(d,)) for d in range(2)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    torch.cuda.set_device(0)
    full = engines[0].generate(prompt, 6, sp).cpu().numpy()
    assert np.array_equal(out[0], full[:, :150])
    agree = (out[1] == full[:, 150:]).mean()       # other GPU, same arithmetic: identical up to bf16 near-ties
    assert agree > (0.999 if precision == 'fp32' else 0.98)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_continuous_batcher_over_two_replicas(monkeypatch):
    """One ContinuousBatcher, two model replicas (one per GPU) pulling from the same queue: every request still gets
    the words it gets alone (greedy), whichever replica served it."""
    monkeypatch.setenv('AWD_PRECISION', 'fp32')
    from pytorch_models import awd_lstm
    from awd_b200.scheduler import ContinuousBatcher
    from awd_b200 import SamplingParams
    from awd_b200.text import format_tokens, exclude_token_ids
    cfg = on.LMConfig(300, 16, 24, 3)
    sd = on.init_state_dict(cfg, 3)
    for k in sd:
        if 'weight_hh' in k or 'weight_ih' in k:
            sd[k] = sd[k] * 4.0
    sd['0.encoder.weight'] = sd['0.encoder_dp.emb.weight'] = sd['1.decoder.weight'] = sd['1.decoder.weight'] * 10.0
    models = []
    for d in range(2):
        m = awd_lstm.get_language_model(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, pad_token=1).cuda(d)
        m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in sd.items()})
        m.eval()
        m.itos = on.synthetic_itos(cfg.vocab_sz)
        m.stoi = {w: i for i, w in enumerate(m.itos)}
        models.append(m)
    m0 = models[0]
    reqs = [([m0.itos[10 + (7 * i) % 200] for _ in range(1 + i % 3)], 2 + (5 * i) % 13) for i in range(40)]
    sp = SamplingParams(greedy=True, banned=exclude_token_ids(m0.stoi))
    torch.cuda.set_device(0)
    want = []
    for words, l in reqs:
        ids = m0.generate_ids(torch.tensor([m0.stoi[w] for w in words]).view(-1, 1), l, sp).cpu()[:, 0].tolist()
        want.append(format_tokens(words, [m0.itos[i] for i in ids]))
    cb = ContinuousBatcher(models, slots=4, chunk=4, greedy=True)
    texts = [f.result(timeout=120) for f in [cb.submit_async(w, l) for w, l in reqs]]
    cb.close()
    assert texts == want
