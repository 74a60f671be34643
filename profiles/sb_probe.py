#!/usr/bin/env python
"""Small-batch operating points of the strong split of BASELINE config 3 (1024 streams over N GPUs):
tokens/s, us per token and the in-kernel phase breakdown at PROBE_BS streams (default 128,256) on one GPU."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))
import torch
from awd_b200 import Engine, SamplingParams
from bench import weights_numpy, V, E, H, L
eng = Engine(V, E, H, L, precision='bf16'); eng.load_state_dict(weights_numpy(0))
T = int(os.environ.get('PROBE_T', 256))
greedy = bool(os.environ.get('PROBE_GREEDY'))
for bs in [int(x) for x in os.environ.get('PROBE_BS', '128,256').split(',')]:
    p = torch.randint(6, V, (1, bs)).cuda(); o = torch.empty(T, bs, dtype=torch.int64, device='cuda')
    sp = lambda i: SamplingParams(banned=[3, 4, 5], seed=10 + i, greedy=greedy)
    for i in range(3): eng.generate(p, T, sp(i), out=o, sync=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    eng.profile(True)
    e0.record()
    for i in range(n): eng.generate(p, T, sp(5 + i), out=o, sync=False)
    e1.record(); torch.cuda.synchronize()
    prof = eng.profile_read(); eng.profile(False)
    ms = e0.elapsed_time(e1) / n
    ph = {k: round(v[0] * 1e3 / max(v[1], 1), 2) for k, v in prof.items() if v[1]}
    hbm = (2 * 44271600 + bs * (43200 + 800 + 8)) / 6554.6e9
    print(f'bs {bs}: {bs * T / ms / 1e3:.3f} M tok/s, {ms * 1e3 / T:.2f} us/token (HBM bound {hbm * 1e6:.1f} us -> frac {hbm / (ms * 1e-3 / T):.3f}) phases us/token {ph}', flush=True)
