"""Serving-side throughput: N requests through BatchScheduler (the object behind GET /inference) on one GPU,
ids -> text included.  Requests are the reference's own shape (1-word prompt, l new words)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))
os.environ.setdefault('AWD_PRECISION', 'bf16')
import torch
from pytorch_models import awd_lstm
from awd_b200.scheduler import BatchScheduler
from bench import weights_numpy, V, E, H, L
N = int(os.environ.get('SERVE_N', 8192)); LNEW = int(os.environ.get('SERVE_L', 256))
m = awd_lstm.get_language_model(V, E, H, L, pad_token=1).cuda()
m.load_state_dict({k: torch.from_numpy(v) for k, v in weights_numpy(0).items()})
m.eval()
m.itos = ['xxunk', 'xxpad', 'xxbos', 'xxfld', 'xxup', 'xxrep', '.', ',', '!', "'s"] + ['w%d' % i for i in range(V - 10)]
m.stoi = {w: i for i, w in enumerate(m.itos)}
import random
random.seed(0)
MIXED = os.environ.get('SERVE_MIXED') == '1'            # request lengths uniform in [16, LNEW] instead of all LNEW
lens = [random.randint(16, LNEW) if MIXED else LNEW for _ in range(N)]
if os.environ.get('SERVE_MODE') == 'continuous':
    from awd_b200.scheduler import ContinuousBatcher
    s = ContinuousBatcher(m, slots=int(os.environ.get('SERVE_SLOTS', 1024)), chunk=int(os.environ.get('SERVE_CHUNK', 32)))
    name = f'ContinuousBatcher(slots={s.slots}, chunk={s.chunk})'
else:
    s = BatchScheduler(m, max_batch=1024, max_wait_ms=5.0)
    name = 'BatchScheduler(max_batch=1024)'
for f in [s.submit_async(['w%d' % (i + 100)], 64) for i in range(1024)]: f.result()      # warm-up
t0 = time.perf_counter()
futs = [s.submit_async(['w%d' % (i % 50000 + 100)], lens[i]) for i in range(N)]
texts = [f.result() for f in futs]
dt = time.perf_counter() - t0
s.close()
words = sum(lens)
print(f'{name}: {N} requests, {words} words ({"mixed" if MIXED else "equal"} lengths) in {dt * 1e3:.0f} ms -> {words / dt / 1e6:.2f} M words/s, '
      f'{N / dt:.0f} requests/s; stats {s.stats}; sample: {texts[0][:50]!r}')
