import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/pytorch-zappa-serverless_b200')
import torch
from awd_b200 import Engine, SamplingParams
from bench import weights_numpy, V, E, H, L
eng = Engine(V, E, H, L, precision='bf16'); eng.load_state_dict(weights_numpy(0))
sp = SamplingParams(banned=[3,4,5], seed=1)
for bs in (8, 1):
    p = torch.randint(6, V, (1, bs)).cuda(); o = torch.empty(119, bs, dtype=torch.int64, device='cuda')
    eng.generate(p, 119, sp, out=o); eng.generate(p, 119, sp, out=o); torch.cuda.synchronize()
