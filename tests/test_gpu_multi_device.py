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
