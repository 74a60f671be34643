"""Smallest run of every schedule of the persistent generation kernel (one, two, three, six and eight M tiles) and of the few-stream
kernel: a quick smoke run, and the command compute-sanitizer would wrap (`compute-sanitizer --tool memcheck python
profiles/sanitize_small.py`; on this pool compute-sanitizer is closed, so memory safety rests on the bounded spins that trap,
the parity tests at every schedule and the oracle comparisons)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))
import numpy as np, torch
from awd_b200 import Engine, SamplingParams
from bench import weights_numpy
V = 3000
eng = Engine(V, 400, 1150, 3, precision='bf16')
eng.load_state_dict(weights_numpy(0, V=V))
for bs in [int(x) for x in os.environ.get('SAN_BS', '3,100,200,300,700,1024').split(',')]:
    p = torch.randint(6, V, (2, bs)).cuda()
    ids = eng.generate(p, 3, SamplingParams(banned=[3, 4, 5], seed=bs)).cpu().numpy()
    assert ids.min() >= 1 and ids.max() < V
    print('bs', bs, 'ok', ids[:, 0], flush=True)
