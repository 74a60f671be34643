"""Times the stand-alone decoder + sampling-race kernel (awd_decode_race_shard over the whole vocabulary)
on 1024 streams."""
import sys, os, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))
import torch
from awd_b200 import Engine, SamplingParams, _lib
from bench import weights_numpy, V, E, H, L
bs = int(os.environ.get('PROBE_BS', 1024))
eng = Engine(V, E, H, L, precision='bf16'); eng.load_state_dict(weights_numpy(0))
lib = _lib.load()
p = torch.randint(6, V, (2, bs)).cuda(); o = torch.empty(1, bs, dtype=torch.int64, device='cuda')
keys = torch.empty(bs, device='cuda'); ids = torch.empty(bs, dtype=torch.int64, device='cuda'); ban = torch.empty(bs, 64, device='cuda')
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
ptr = lambda t: C.c_void_p(t.data_ptr())
for greedy in (False, True):
    sp = SamplingParams(banned=[3, 4, 5], seed=1, greedy=greedy)
    eng.generate(p, 1, sp, out=o)
    s, _keep = sp.to_c()
    def run(i):
        _lib.check(lib.awd_decode_race_shard(eng._h, 0, V, C.byref(s), i, ptr(keys), ptr(ids), ptr(ban), st))
    for i in range(5): run(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 100
    e0.record()
    for i in range(n): run(i)
    e1.record(); torch.cuda.synchronize()
    print(f"greedy {greedy}: {e0.elapsed_time(e1) * 1e3 / n:.1f} us per decode+race (incl. shard_partial kernel)", flush=True)
