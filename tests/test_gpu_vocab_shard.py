"""Vocabulary-sharded decoder (BASELINE config 4).  The race keys are functions of
(chunk, stream, step, seed) only, so ANY sharding must reproduce the unsharded engine's ids exactly --
checked here with logical shards on one GPU, and over NCCL when two GPUs are present."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

from oracle import awd_numpy as on

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _model(V=5000, seed=51):
    cfg = on.LMConfig(V, 400, 1150, 3)
    sd = on.init_state_dict(cfg, seed, decoder_bias_scale=0.3)
    sd['1.decoder.bias'][:40] += 7.0        # disallowed low ids lead, so that the fallback rule is exercised
    return cfg, sd


@pytest.mark.parametrize('n_shards', [1, 2, 3, 8])
def test_logical_shards_reproduce_unsharded_ids(n_shards):
    from awd_b200 import Engine, SamplingParams
    from awd_b200.vocab_shard import generate_vocab_sharded, shard_bounds
    cfg, sd = _model()
    e = Engine(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, precision='bf16')
    e.load_state_dict(sd)
    prompt = torch.randint(6, 5000, (2, 40), generator=torch.Generator().manual_seed(1)).cuda()
    b = shard_bounds(5000, n_shards)
    assert b[0][0] == 0 and b[-1][1] == 5000 and all(x[1] == y[0] and y[0] % 256 == 0 for x, y in zip(b, b[1:]))
    for sp in [SamplingParams(greedy=True, banned=[3, 4, 5]), SamplingParams(banned=[3, 4, 5], seed=9),
               SamplingParams(n_draws=2, banned=list(range(1, 30)), seed=10, temperature=0.8)]:
        want = e.generate(prompt, 6, sp).cpu().numpy()                       # unsharded persistent kernel
        got = generate_vocab_sharded(e, prompt, 6, sp, logical_shards=n_shards).cpu().numpy()
        assert np.array_equal(got, want), (n_shards, sp, (got != want).mean())
        if sp.n_draws == 2:
            assert np.isin(got, [0] + list(range(1, 30))).mean() > 0.2       # the fallback branch really fired
        if n_shards == 3:       # the same token loop captured once as a CUDA graph and replayed (twice: a replay must be repeatable)
            from awd_b200.vocab_shard import VocabShardedGenerator
            gen = VocabShardedGenerator(e, prompt.shape[1], logical_shards=n_shards, use_graph=True)
            for _ in range(2):
                got = gen.generate(prompt, 6, sp).cpu().numpy()
                assert np.array_equal(got, want), ('graph', sp)
            e.reset(prompt.shape[1])


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _nccl_worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))
    import torch.distributed as dist
    from awd_b200 import Engine, SamplingParams
    from awd_b200.vocab_shard import generate_vocab_sharded
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
    cfg, sd = _model()
    e = Engine(cfg.vocab_sz, cfg.emb_sz, cfg.n_hid, cfg.n_layers, precision='bf16', device=rank)
    e.load_state_dict(sd)
    prompt = torch.randint(6, 5000, (2, 40), generator=torch.Generator().manual_seed(1)).cuda()
    sp = SamplingParams(banned=[3, 4, 5], seed=9)
    got = generate_vocab_sharded(e, prompt, 6, sp).cpu().numpy()
    want = e.generate(prompt, 6, sp).cpu().numpy()
    ok = np.array_equal(got, want)
    if os.environ.get('AWD_TEST_NCCL_GRAPH') == '1':                      # NCCL all-gathers inside a CUDA graph (opt-in)
        from awd_b200.vocab_shard import VocabShardedGenerator
        gen = VocabShardedGenerator(e, prompt.shape[1], use_graph=True)
        for _ in range(2):
            ok = ok and np.array_equal(gen.generate(prompt, 6, sp).cpu().numpy(), want)
    flag = torch.tensor([int(ok)], device='cuda')
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        open(out, 'w').write('ok' if int(flag) == 1 else 'mismatch')
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_two_ranks_over_nccl_reproduce_unsharded_ids(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / 'r.txt')
    mp.spawn(_nccl_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert open(out).read() == 'ok'
