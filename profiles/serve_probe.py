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
s = BatchScheduler(m, max_batch=1024, max_wait_ms=5.0)
for f in [s.submit_async(['w%d' % (i + 100)], LNEW) for i in range(1024)]: f.result()      # warm-up batch
t0 = time.perf_counter()
futs = [s.submit_async(['w%d' % (i % 50000 + 100)], LNEW) for i in range(N)]
texts = [f.result() for f in futs]
dt = time.perf_counter() - t0
s.close()
print(f'{N} requests x {LNEW} words in {dt * 1e3:.0f} ms -> {N * LNEW / dt / 1e6:.2f} M words/s, {N / dt:.0f} requests/s; '
      f'scheduler stats {s.stats}; sample: {texts[0][:60]!r}')
