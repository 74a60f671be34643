import sys, os, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/pytorch-zappa-serverless_b200')
import torch
from awd_b200 import Engine, SamplingParams
from bench import weights_numpy, V, E, H, L
eng = Engine(V, E, H, L, precision='bf16'); eng.load_state_dict(weights_numpy(0))
sp = SamplingParams(banned=[3,4,5], seed=1)
def t(P, n, bs=1, reps=10):
    p = torch.randint(6, V, (P, bs)).cuda(); o = torch.empty(n, bs, dtype=torch.int64, device='cuda')
    for _ in range(3): eng.generate(p, n, sp, out=o)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): eng.generate(p, n, sp, out=o)
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e6
a = t(120, 0); b = t(1, 0); c = t(20, 100); d = t(1, 119)
print(f'prompt-only 120 steps: {a:.0f} us -> {(a-b)/119:.1f} us/step (layers only); launch+setup {b:.0f} us')
print(f'20+100: {c:.0f} us; 1+119: {d:.0f} us -> decode step {(d-b)/119:.1f} us -> decoder+sample part {(d-a)/119:.1f} us')
for bs in (2, 4, 8): print('bs', bs, f'{t(1, 119, bs):.0f} us')
