"""World-size-2 checks of the multi-GPU plan on CPU (gloo): streams are sharded by batch with
replicated weights and NO data-path collective.  Each rank runs the oracle on its own
shard; rank 0 gathers and compares with the unsharded run -- the property the engine's shards must
(and, in tests/test_gpu_parity.py, do) satisfy.  Also covers the max-over-ranks timing reduction
bench.py uses."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    from oracle import awd_numpy as on
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    cfg = on.LMConfig(90, 8, 12, 2)
    sd = on.init_state_dict(cfg, 3)
    for k in sd:
        if 'weight_ih' in k or 'weight_hh' in k:
            sd[k] = sd[k] * 6.0
    p = on.effective_params(sd, cfg)
    B = 8
    prompt = np.random.default_rng(0).integers(6, 90, (2, B))
    lo, hi = rank * B // world, (rank + 1) * B // world
    ids, _ = on.generate_ids(p, cfg, prompt[:, lo:hi], 6, [3, 4, 5])            # this rank's streams only
    mine = torch.from_numpy(ids)
    gathered = [torch.zeros_like(mine) for _ in range(world)] if rank == 0 else None
    dist.gather(mine, gathered, dst=0)
    t = torch.tensor([10.0 + rank], dtype=torch.float64)                        # bench.py: MAX over ranks
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        full, _ = on.generate_ids(p, cfg, prompt, 6, [3, 4, 5])
        ok = np.array_equal(torch.cat(gathered, 1).numpy(), full) and float(t) == 10.0 + world - 1
        open(out, 'w').write('ok' if ok else 'mismatch')
    dist.barrier()
    dist.destroy_process_group()


def test_batch_sharding_reproduces_unsharded_run_world2(tmp_path):
    out = str(tmp_path / 'result.txt')
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert open(out).read() == 'ok'


def test_scheduler_spreads_batches_over_replicas():
    sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))
    from awd_b200.scheduler import BatchScheduler
    from test_host_cpu import _FakeModel
    m0, m1 = _FakeModel(), _FakeModel()
    s = BatchScheduler([m0, m1], max_batch=2, max_wait_ms=20)
    futs = [s.submit_async([f'w{10 + i}'], 4) for i in range(12)]
    texts = [f.result(timeout=10) for f in futs]
    s.close()
    assert len(texts) == 12 and all(t.startswith(' w1') or t.startswith(' w2') for t in texts)
    assert len(m0.calls) + len(m1.calls) >= 6 and m0.calls and m1.calls      # both replicas served batches
