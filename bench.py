#!/usr/bin/env python
"""bench.py -- generated tokens/sec of the AWD-LSTM hot path on B200 (BASELINE.json metric).

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): AWD-LSTM as
shipped (emb 400, hidden 1150, 3 layers, vocab 60k, random-init), 1024 concurrent streams x 256
new tokens from 1-token prompts, sampling with the reference rule (main.py:61-68).  One "step" is
one whole generation call.  With N GPUs the 1024 streams are SPLIT over the ranks, 1024/N per GPU on
a weight replica, no collective on the data path (BASELINE.json: "sharded by batch across 1/2/4/8
B200"): strong scaling.  The weak point (1024 streams on every GPU) is an extra key.

    python bench.py [--gpus N --steps K --warmup W] [--precision fp32|bf16] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = tokens/s with prompts resident in HBM and ids left in
HBM; `e2e` = the same call through the C ABI with HOST (pinned) prompt and output buffers, copies
inside the timed region.  `--impl reference` times the CPU port of the reference's operator
sequence (oracle/awd_torch_cpu.py) on the host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))

import numpy as np
import torch

V, E, H, L = 60000, 400, 1150, 3
STREAMS, NEW_TOKENS, PROMPT = 1024, 256, 1
METRIC = 'generated tokens/sec (whole box) AWD-LSTM 3x1150'
BANNED = [3, 4, 5]            # xxfld, xxup, xxrep of the synthetic vocabulary

# algorithmic work per generated token
FLOP_PER_TOKEN = 2 * (8_970_000 + 11_220_000 + 24_000_000)
DEC_FLOP_PER_TOKEN = 2 * 24_000_000
WEIGHT_ELEMS = 20_190_000 + 21_600 + 24_000_000 + 60_000


def algorithmic_bytes_per_step(B: int, s: int) -> int:
    """Bytes one generation step must move per GPU: weights once + state read/write + one
    gathered row + token in/out."""
    return s * WEIGHT_ELEMS + B * (2 * 5400 * 4 + s * E + 8)


# per phase class of one token step (bf16: s = 2): weights of the phase once + the state it reads and writes
DEC_WEIGHT_ELEMS = 24_000_000 + 60_000
LSTM_WEIGHT_ELEMS = 20_190_000 + 21_600
LSTM_FLOP_PER_TOKEN = 2 * (8_970_000 + 11_220_000)


def phase_bytes(B: int, s: int):
    return {'decoder': s * DEC_WEIGHT_ELEMS + B * (s * E + 8), 'lstm': s * LSTM_WEIGHT_ELEMS + B * (2 * 5400 * 4 + s * E)}


def measured_traffic(streams: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of mega_generate_kernel per token step, from the committed ncu
    capture of this round (profiles/r02_mega_ncu_traffic.json, written by profiles/ncu_traffic.py from the raw csv).
    None when no capture exists for this operating point: a constant from an older kernel is not evidence."""
    path = os.path.join(ROOT, 'profiles', 'r02_mega_ncu_traffic.json')
    if not os.path.exists(path):
        return None, 'no committed ncu capture'
    d = json.load(open(path))
    e = d.get(str(streams))
    if not e:
        return None, f'no committed ncu capture at {streams} streams'
    return e['dram_bytes_per_token'], e.get('source', path)


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return d['hbm_gbs'], d['bf16_tflops'], d.get('bf16_tflops_sustained', d['bf16_tflops']), 'measured'
    return 6650.0, 1590.0, 1400.0, 'fallback'


def weights_numpy(seed=0, V=V, E=E, H=H, L=L):
    """Random-init with the reference constructor's distributions (pure numpy here so that the
    product arm does not import oracle/)."""
    rng = np.random.default_rng(seed)
    sd = {}
    enc = rng.uniform(-0.1, 0.1, (V, E)).astype(np.float32)
    sd['0.encoder.weight'] = sd['0.encoder_dp.emb.weight'] = sd['1.decoder.weight'] = enc
    for l in range(L):
        i, h = (E if l == 0 else H), (H if l != L - 1 else E)
        k = 1.0 / np.sqrt(h)
        p = f'0.rnns.{l}.'
        whh = rng.uniform(-k, k, (4 * h, h)).astype(np.float32)
        sd[p + 'weight_hh_l0_raw'] = sd[p + 'module.weight_hh_l0'] = whh
        sd[p + 'module.weight_ih_l0'] = rng.uniform(-k, k, (4 * h, i)).astype(np.float32)
        sd[p + 'module.bias_ih_l0'] = rng.uniform(-k, k, (4 * h,)).astype(np.float32)
        sd[p + 'module.bias_hh_l0'] = rng.uniform(-k, k, (4 * h,)).astype(np.float32)
    sd['1.decoder.bias'] = np.zeros(V, np.float32)
    order = ['0.encoder.weight', '0.encoder_dp.emb.weight']
    for l in range(L):
        p = f'0.rnns.{l}.'
        order += [p + 'weight_hh_l0_raw', p + 'module.weight_ih_l0', p + 'module.weight_hh_l0',
                  p + 'module.bias_ih_l0', p + 'module.bias_hh_l0']
    order += ['1.decoder.weight', '1.decoder.bias']
    return {k: sd[k] for k in order}


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for r in self.rows if len(r) == 7]
        if not rows:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['no samples']}
        sm = sorted(float(r[0]) for r in rows)
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any(r[3 + i] == 'Active' for r in rows)]
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': float(rows[0][1]), 'reasons': reasons,
                'power_w_max': max(float(r[2]) for r in rows), 'samples': len(rows)}


def flush_l2(buf):
    buf.add_(1)       # 256 MiB read+write, larger than the 126 MB L2


def route_latency(dev, local_rank, precision, requests=100):
    """C1 as BASELINE.json words it -- "via the Flask predict route": checkpoint + vocabulary files in the local cache,
    main.load_lm_and_predict() behind GET /inference (the WSGI callable main.app, called in-process by one client), the
    request queue, the batch scheduler with AWD_MAX_WAIT_MS=0 and the formatter thread all inside the timed call."""
    import tempfile
    import pickle
    from urllib.parse import quote
    tmp = tempfile.mkdtemp(prefix='awd_bench_')
    root = os.path.join(tmp, 'models', 'rjokes')
    os.makedirs(root)
    torch.save({k: torch.from_numpy(v) for k, v in weights_numpy(0).items()}, os.path.join(root, 'rjokes.model.pth'))
    head = ['xxunk', 'xxpad', 'xxbos', 'xxfld', 'xxup', 'xxrep', '.', ',', '!', '\n', "'s", "n't", '?']
    itos = head + [f'w{i}' for i in range(len(head), V)]
    with open(os.path.join(root, 'rjokes.itos.pkl'), 'wb') as f:
        pickle.dump(itos, f)
    os.environ.update({'AWD_CACHE_DIR': tmp, 'models_bucket': 'bench', 'AWD_PRECISION': precision, 'AWD_MAX_WAIT_MS': '0',
                       'AWD_DEVICE': str(local_rank)})
    import main as product_main                      # pytorch-zappa-serverless_b200/main.py (bench.py itself runs as __main__)
    g = torch.Generator().manual_seed(77)
    lat = []
    try:
        for i in range(requests + 3):
            words = [itos[int(t)] for t in torch.randint(13, V, (20,), generator=g)]
            query = 'l=100&prompt=' + quote(' '.join(words))
            t0 = time.perf_counter()
            if hasattr(product_main.app, 'test_client'):
                body = product_main.app.test_client().get('/inference?' + query).data
            else:
                body = b''.join(product_main.app({'PATH_INFO': '/inference', 'REQUEST_METHOD': 'GET', 'QUERY_STRING': query}, lambda s_, h_: None))
            lat.append((time.perf_counter() - t0) * 1e3)
            text = json.loads(body)['response']['text']
            assert len(text.split(' ')) > 100
    finally:
        for sch in (product_main._SCHED, product_main._CONT):
            if sch is not None:
                sch.close()
        import shutil
        shutil.rmtree(tmp, ignore_errors=True)
    lat = sorted(lat[3:])
    return {'p50_ms': lat[len(lat) // 2], 'p90_ms': lat[(len(lat) * 9) // 10], 'requests': len(lat),
            'workload': 'C1 through GET /inference (main.app, in-process WSGI call, one client): 20-word prompt + 100 words, model loaded '
                        'from checkpoint files, BatchScheduler with AWD_MAX_WAIT_MS=0, text formatting included',
            'tokens_per_s': 100 / (lat[len(lat) // 2] / 1e3)}


def bench_c4(args, rank, world, local_rank):
    """BASELINE.json configs[3]: the WikiText-103 vocabulary (267 735 rows) with the decoder rows sharded over the ranks, 1024
    replicated streams.  Per token: fused decoder + race over the rank's rows, ONE NCCL all-gather of 68 floats per stream and
    rank, merge, one replicated LSTM step; the whole token loop is one CUDA graph.  Equality with the unsharded engine is
    checked in this same job, on every rank."""
    import ctypes as C
    import torch.distributed as dist
    from awd_b200 import Engine, SamplingParams, _lib
    from awd_b200.vocab_shard import VocabShardedGenerator, shard_bounds, BLOCK_FLOATS_PER_STREAM
    V4, B, T = 267735, args.streams, (args.new_tokens if args.new_tokens != NEW_TOKENS else 64)
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    eng = Engine(V4, E, H, L, precision='bf16', device=local_rank)
    eng.load_state_dict(weights_numpy(0, V=V4))
    stream = torch.cuda.current_stream(dev)
    prompt_host = torch.randint(6, V4, (PROMPT, B), generator=torch.Generator().manual_seed(1)).pin_memory()
    prompt = prompt_host.to(dev)
    out = torch.empty(T, B, dtype=torch.int64, device=dev)
    out_host = torch.empty(T, B, dtype=torch.int64).pin_memory()
    sp = SamplingParams(banned=BANNED, seed=1234)
    # ---- parity inside the job: sharded == unsharded, sampled and greedy, every rank
    ok = 1
    for s_ in (SamplingParams(banned=BANNED, seed=7), SamplingParams(banned=BANNED, greedy=True)):
        want = eng.generate(prompt, 8, s_).clone()
        g0 = VocabShardedGenerator(eng, B, logical_shards=None if world > 1 else 4)
        got = g0.generate(prompt, 8, s_)
        ok = ok and int(torch.equal(want, got))
    flag = torch.tensor([ok], device=dev)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ids_equal = bool(int(flag.item()))
    use_graph = world == 1 or os.environ.get('AWD_C4_GRAPH') == '1'      # NCCL inside a CUDA graph: opt-in
    gen = VocabShardedGenerator(eng, B, logical_shards=None if world > 1 else 1, use_graph=use_graph)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    def timed(n, host):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
        barrier()
        for i in range(n):
            ev[i][0].record(stream)
            if host:
                prompt.copy_(prompt_host, non_blocking=True)
            gen.generate(prompt, T, sp, out=out)
            if host:
                out_host.copy_(out, non_blocking=True)
            ev[i][1].record(stream)
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in ev)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(max(args.warmup, 1)):
        gen.generate(prompt, T, sp, out=out)
    clocks = ClockSampler(local_rank)
    clocks.start()
    l0 = eng.launch_count()
    ms_dev = timed(args.steps, False)
    launches = eng.launch_count() - l0          # (graph replays launch without passing through the library's counter: see note)
    ms_e2e = timed(args.steps, True)
    clk = clocks.stop()
    # ---- the pieces: the rank's decoder + race alone, and the all-gather alone
    v0, v1 = shard_bounds(V4, world)[rank] if world > 1 else (0, V4)
    s_c, keep = sp.to_c()
    st = C.c_void_p(stream.cuda_stream)
    n_rep = 50
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eng.reset(B); eng.forward(prompt, need_logits=False, need_raw=False)
    for i in range(n_rep + 5):
        if i == 5:
            e0.record(stream)
        _lib.check(eng.lib.awd_decode_race_shard_packed(eng._h, v0, v1, C.byref(s_c), i, C.c_void_p(gen.blocks[rank if world > 1 else 0].data_ptr()), st))
    e1.record(stream); torch.cuda.synchronize(dev)
    dec_us = e0.elapsed_time(e1) * 1e3 / n_rep
    ag_us = None
    if world > 1:
        for i in range(n_rep + 5):
            if i == 5:
                e0.record(stream)
            dist.all_gather_into_tensor(gen.blocks.view(-1), gen.blocks[rank])
        e1.record(stream); torch.cuda.synchronize(dev)
        ag_us = e0.elapsed_time(e1) * 1e3 / n_rep
    # the same workload without sharding, on this one GPU (persistent kernel, whole vocabulary)
    for _ in range(2):
        eng.generate(prompt, T, sp, out=out)
    e0.record(stream)
    for _ in range(2):
        eng.generate(prompt, T, sp, out=out)
    e1.record(stream); torch.cuda.synchronize(dev)
    unsharded_tok_s = 2 * B * T / (e0.elapsed_time(e1) / 1e3)
    hbm, tf_burst, tf_sus, which = peaks()
    tokens = args.steps * B * T
    value = tokens / (ms_dev / 1e3)
    dec_flop = 2.0 * B * (v1 - v0) * E
    line = {
        'metric': METRIC + ', vocabulary 267735 with the decoder sharded over the GPUs (BASELINE config 4)', 'value': value, 'unit': 'tokens/s',
        'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms_dev / args.steps, 'higher_is_better': True,
        'scaling': 'strong', 'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
        'config': {'workload': f'C4: vocab {V4} (decoder rows sharded {world} ways, {v1 - v0} rows on rank {rank}), {B} replicated streams x {T} new tokens, '
                               f'emb {E} hid {H} layers {L}, reference sampling rule; token loop ' + ('= one CUDA graph per rank' if use_graph else 'enqueued from the host without any synchronisation') + ' (decoder+race, 1 NCCL all-gather, merge, LSTM step)',
                   'parallelism': f'vocabulary-sharded decoder x{world}, recurrence replicated, one all-gather of {B * BLOCK_FLOATS_PER_STREAM * 4} bytes per rank and token'},
        'clocks': clk,
        'e2e': {'value': tokens / (ms_e2e / 1e3), 'unit': 'tokens/s', 'h2d_bytes_per_step': prompt_host.numel() * 8, 'd2h_bytes_per_step': out_host.numel() * 8,
                'ms_per_step': ms_e2e / args.steps},
        'gpu_launches': (launches + (2 * args.steps * T * 8 if use_graph else 0)) * world,
        'gpu_launches_note': 'per token and rank: row-bound memset node aside, decoder+race, shard partial, merge, token copy, gather + 3 LSTM layers = 8 kernels inside a replayed CUDA graph (counted from the graph: replays do not pass through the library counter), both timed regions, all ranks',
        'roofline': {'kernel': f'tc_gemm_kernel<256, TcEpiSample> over this rank\'s {v1 - v0} decoder rows + shard_partial_kernel, M={B}', 'bound': 'tensor',
                     'achieved': dec_flop / (dec_us * 1e-6) / 1e12, 'peak': tf_burst, 'unit': 'TFLOP/s', 'frac': dec_flop / (dec_us * 1e-6) / 1e12 / tf_burst,
                     'peak_source': which + ' (burst bf16 cuBLAS: kernel timed alone, back to back, CUDA events)', 'avg_launch_us': dec_us, 'traffic': None,
                     'traffic_source': 'not captured (multi-rank command: no ncu)'},
        'per_token_us': {'whole_step': ms_dev / args.steps * 1e3 / T, 'decoder_race_shard': dec_us, 'all_gather': ag_us},
        'nccl_ranks': world if world > 1 else 0, 'ids_equal_unsharded': ids_equal,
        'unsharded_one_gpu_tokens_per_s': unsharded_tok_s,
    }
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    assert ids_equal, 'vocabulary-sharded ids differ from the unsharded engine'
    return 0


def bench_c5(args, local_rank):
    """BASELINE.json configs[4]: the scaled model (emb 1024, hidden 3072, 4 layers, vocab 60k, bf16): a 4096-token prompt at batch
    1 and batch 8, then 256 generated tokens.  One GPU.  Reported against SURVEY.md 8(d): 559.1 MFLOP per token; 559.4 MB of bf16
    weights; at batch 1 a recurrent step cannot beat streaming W_hh (234.9 MB -> 35.9 us at the measured HBM rate)."""
    from awd_b200 import Engine, SamplingParams
    E5, H5, L5, P5, T5 = 1024, 3072, 4, 4096, 256
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    eng = Engine(V, E5, H5, L5, precision='bf16', device=local_rank)
    eng.load_state_dict(weights_numpy(0, V=V, E=E5, H=H5, L=L5))
    stream = torch.cuda.current_stream(dev)
    hbm, tf_burst, tf_sus, which = peaks()
    w_ih = sum(4 * (H5 if l != L5 - 1 else E5) * (E5 if l == 0 else H5) for l in range(L5))
    w_hh = sum(4 * (H5 if l != L5 - 1 else E5) ** 2 for l in range(L5))
    w_dec = V * E5
    flop_tok = 2 * (w_ih + w_hh + w_dec)
    sp = SamplingParams(banned=BANNED, seed=1234)
    pts = []
    clocks = ClockSampler(local_rank)
    clocks.start()
    l0 = eng.launch_count()
    for bs in (1, 8):
        prompt = torch.randint(6, V, (P5, bs), generator=torch.Generator().manual_seed(bs)).to(dev)
        out = torch.empty(T5, bs, dtype=torch.int64, device=dev)

        def timed(fn, n=3):
            fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(dev)
            e0.record(stream)
            for _ in range(n):
                fn()
            e1.record(stream)
            torch.cuda.synchronize(dev)
            return e0.elapsed_time(e1) / n

        def prefill_forward():
            eng.reset(bs)
            eng.forward(prompt, need_raw=False, last_step_only=True)
        ms_fwd = timed(prefill_forward)                                   # awd_forward: time-batched input projections + wavefront recurrence
        ms_pre = timed(lambda: eng.generate(prompt, 0, sp, out=out[:0], sync=False))        # the prompt inside awd_generate
        ms_all = timed(lambda: eng.generate(prompt, T5, sp, out=out, sync=False))
        step_us = (ms_all - ms_pre) * 1e3 / T5
        bytes_step = 2 * (w_ih + w_hh + w_dec)                             # bf16 weights once per generated token
        pts.append({'streams': bs, 'prefill_ms_awd_forward': ms_fwd, 'prefill_ms_inside_awd_generate': ms_pre,
                    'prefill_tokens_per_s': P5 * bs / (min(ms_fwd, ms_pre) / 1e3),
                    'prefill_bound_ms': {'w_hh_stream_per_step': P5 * 2 * w_hh / (hbm * 1e9) * 1e3, 'tensor': P5 * bs * flop_tok / (tf_sus * 1e12) * 1e3},
                    'generate_us_per_token': step_us, 'generate_hbm_bound_us': bytes_step / (hbm * 1e9) * 1e6,
                    'generate_frac_of_hbm_roofline': bytes_step / (hbm * 1e9) * 1e6 / step_us,
                    'generated_tokens_per_s': bs * 1e6 / step_us})
    launches = eng.launch_count() - l0
    clk = clocks.stop()
    p1 = pts[0]
    line = {
        'metric': METRIC.replace('3x1150', '4x3072 (scaled, BASELINE config 5)'), 'value': pts[1]['generated_tokens_per_s'], 'unit': 'tokens/s', 'n_gpus': 1,
        'steps': 3, 'warmup': 1, 'ms_per_step': None, 'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
        'config': {'workload': f'C5: emb {E5} hid {H5} layers {L5} vocab {V}, bf16; prefill {P5} tokens at batch 1 and 8, then {T5} generated tokens; value = generated tokens/s at batch 8'},
        'clocks': clk, 'gpu_launches': launches,
        'roofline': {'kernel': 'lat_generate_kernel (<= 8 streams: prompt + generation in one persistent launch), generation step at batch 1', 'bound': 'hbm',
                     'achieved': 2 * (w_ih + w_hh + w_dec) / (p1['generate_us_per_token'] * 1e-6) / 1e9, 'peak': hbm, 'unit': 'GB/s',
                     'frac': p1['generate_frac_of_hbm_roofline'], 'peak_source': which, 'traffic': None, 'traffic_source': 'not captured'},
        'points': pts, 'flop_per_token': flop_tok, 'weight_bytes_bf16': 2 * (w_ih + w_hh + w_dec),
    }
    print(json.dumps(line))
    return 0


def run_cpu_port(steps, warmup, new_tokens, streams, threads=None):
    """Times oracle/awd_torch_cpu.py (the reference's operator sequence on torch CPU) -- the only
    place bench.py executes anything under oracle/."""
    from oracle.awd_numpy import LMConfig
    from oracle.awd_torch_cpu import TorchCpuLM
    # torchrun exports OMP_NUM_THREADS=1; the baseline is entitled to every host core it can use
    torch.set_num_threads(threads or len(os.sched_getaffinity(0)))
    lm = TorchCpuLM(weights_numpy(0), LMConfig(V, E, H, L))
    prompt = torch.randint(6, V, (PROMPT, streams), generator=torch.Generator().manual_seed(1))
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            torch.manual_seed(1234 + i)
            lm.bs = streams
            lm.reset()
            z, _ = lm.forward(prompt)                      # prompt step, untimed
            t0 = time.perf_counter()
            for _ in range(new_tokens):                    # main.py:60-71
                ids = lm.generate_step(z, BANNED)
                z, _ = lm.forward(ids[None, :])
            if i >= warmup:
                times.append(time.perf_counter() - t0)
    total = sum(times)
    return steps * streams * new_tokens / total, total / steps * 1e3, torch.get_num_threads()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--precision', default=os.environ.get('AWD_BENCH_PRECISION', 'bf16'), choices=['fp32', 'bf16'])
    ap.add_argument('--streams', type=int, default=STREAMS)
    ap.add_argument('--new-tokens', type=int, default=NEW_TOKENS)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-latency', action='store_true')
    ap.add_argument('--workload', default='c3', choices=['c3', 'c4', 'c5'])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if args.streams % world:
        raise SystemExit(f'--streams {args.streams} does not split over {world} ranks')
    per_gpu = args.streams // world
    config = {'workload': f'C3: {args.streams} streams in total x {args.new_tokens} new tokens, {per_gpu} streams per GPU on {world} GPU(s), '
                          f'{PROMPT}-token prompts, emb {E} hid {H} layers {L} vocab {V}, reference sampling rule (10 draws, banned ids)',
              'streams_total': args.streams, 'streams_per_gpu': per_gpu,
              'parallelism': f'replicas x{world} (streams sharded by batch, weights replicated, no collective)',
              'l2': 'flushed between steps (256 MiB read+write, untimed)'}

    if args.workload == 'c4' and args.impl != 'reference':
        return bench_c4(args, rank, world, local_rank)
    if args.workload == 'c5' and args.impl != 'reference':
        return bench_c5(args, local_rank)
    if args.impl == 'reference':
        if rank != 0:
            return 0
        # bounded sample: every step generates 2 tokens for all streams (about 3-4 s per step)
        sample_tokens = 2
        val, ms, cores = run_cpu_port(args.steps, args.warmup, sample_tokens, args.streams)
        sample = f'{args.streams} streams x {sample_tokens} new tokens per step (after an untimed {PROMPT}-token prompt)'
        print(json.dumps({
            'impl': 'reference', 'metric': METRIC, 'value': val, 'unit': 'tokens/s', 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True,
            'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': config,
            'cpu_baseline': {'value': val, 'unit': 'tokens/s', 'cores': cores, 'kind': 'port', 'sample': sample},
            'e2e': {'value': val, 'unit': 'tokens/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}))
        return 0

    from awd_b200 import Engine, SamplingParams
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    eng = Engine(V, E, H, L, precision=args.precision, device=local_rank)
    eng.load_state_dict(weights_numpy(0))
    B, T = per_gpu, args.new_tokens
    g = torch.Generator().manual_seed(1 + rank)
    prompt_host = torch.randint(6, V, (PROMPT, B), generator=g).pin_memory()
    out_host = torch.empty(T, B, dtype=torch.int64).pin_memory()
    prompt_dev = prompt_host.to(dev)
    out_dev = torch.empty(T, B, dtype=torch.int64, device=dev)
    flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    greedy = bool(os.environ.get('AWD_BENCH_GREEDY'))     # experiment knob: argmax instead of the reference rule
    sp = lambda i: SamplingParams(banned=BANNED, seed=1234 + 1000 * rank + i, greedy=greedy)
    stream = torch.cuda.current_stream(dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    def timed(prompt, out, n, profile):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
        barrier()
        if profile:
            eng.profile(True)
        for i in range(n):
            flush_l2(flush)
            ev[i][0].record(stream)
            eng.generate(prompt, T, sp(i), out=out, sync=False)
            ev[i][1].record(stream)
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in ev)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for i in range(args.warmup):
        eng.generate(prompt_dev, T, sp(100 + i), out=out_dev, sync=False)
    eng.check_errors()
    clocks = ClockSampler(local_rank)
    clocks.start()
    l0 = eng.launch_count()
    ms_dev = timed(prompt_dev, out_dev, args.steps, profile=True)
    launches = eng.launch_count() - l0
    prof = eng.profile_read()
    eng.profile(False)
    ms_e2e = timed(prompt_host, out_host, args.steps, profile=False)
    clk = clocks.stop()
    eng.check_errors()
    assert int(out_host.min()) >= 1 and int(out_host.max()) < V and not set(BANNED) & set(out_host.unique().tolist())

    tokens = args.steps * B * T * world          # = steps x streams_total x new tokens
    value = tokens / (ms_dev / 1e3)
    e2e = tokens / (ms_e2e / 1e3)
    if world > 1:
        t = torch.tensor([launches], device=dev, dtype=torch.int64)
        dist.all_reduce(t)
        launches = int(t.item())

    hbm, tf_burst, tf_sus, which = peaks()
    s_bytes = 2 if args.precision == 'bf16' else 4
    mega = args.precision == 'bf16' and os.environ.get('AWD_MEGA', '1') != '0'
    # which roofline binds a token step at B streams per GPU: min(HBM, tensor / FMA) in tokens/s
    step_bytes = algorithmic_bytes_per_step(B, s_bytes)
    hbm_tok_s = B / (step_bytes / (hbm * 1e9))
    flop_peak = tf_sus if args.precision == 'bf16' else 74.4
    tensor_tok_s = flop_peak * 1e12 / FLOP_PER_TOKEN
    bound = 'hbm' if hbm_tok_s <= tensor_tok_s else 'tensor'
    per_gpu_tok_s = value / world

    def entry(name, seconds_per_token, flop, nbytes):
        """One phase class against the roofline that binds it at this batch: achieved = algorithmic work / measured time."""
        if not seconds_per_token:
            return None
        t_hbm, t_flop = nbytes / (hbm * 1e9), flop * B / (flop_peak * 1e12)
        if t_hbm >= t_flop:
            ach, peak, unit, bnd = nbytes / seconds_per_token / 1e9, hbm, 'GB/s', 'hbm'
        else:
            ach, peak, unit, bnd = flop * B / seconds_per_token / 1e12, flop_peak, 'TFLOP/s', 'tensor'
        return {'phase': name, 'bound': bnd, 'achieved': ach, 'peak': peak, 'unit': unit, 'frac': ach / peak,
                'us_per_token': seconds_per_token * 1e6, 'algorithmic_bytes': nbytes, 'algorithmic_flop': flop * B}

    us = {k: (v[0] * 1e3 / max(v[1], 1)) for k, v in prof.items() if v[1]}           # per launch / per token and phase
    pb = phase_bytes(B, s_bytes)
    dec_s = us.get('decoder', 0.0) * 1e-6
    lstm_s = (us.get('step', 0.0) * L + (us.get('sample', 0.0) if mega else 0.0)) * 1e-6   # the gather+finalize phase feeds layer 0
    step_s = 1.0 / (per_gpu_tok_s / B)                                                  # seconds per token step, all phases, measured end to end
    phases = [entry('decoder + sampling race', dec_s, DEC_FLOP_PER_TOKEN, pb['decoder']),
              entry('LSTM layers (+ finalize/gather)', lstm_s, LSTM_FLOP_PER_TOKEN, pb['lstm']),
              entry('whole token step', step_s, FLOP_PER_TOKEN, step_bytes)]
    whole = phases[2]
    traffic, traffic_src = measured_traffic(B) if mega else (None, 'not captured for this path')
    roofline = {
        'kernel': ('mega_generate_kernel, whole token step (persistent: decoder+race, finalize+gather, 3 LSTM layers per token'
                   if mega else 'per-kernel path, whole token step (') + f', M={B} streams per GPU)',
        'timing': ('whole step: CUDA events around the generation call on the launching stream; phases: in-kernel %globaltimer '
                   'around each phase (CTA 0), summed over the timed region -- the persistent kernel has no per-phase launch to bracket'
                   if mega else 'CUDA events recorded around every launch on the launching stream'),
        'bound': whole['bound'], 'achieved': whole['achieved'], 'peak': whole['peak'], 'unit': whole['unit'], 'frac': whole['frac'],
        'peak_source': which + (' (hbm_gbs: copy bandwidth)' if whole['bound'] == 'hbm' else ' (sustained bf16 cuBLAS)'),
        'traffic': traffic, 'traffic_unit': 'bytes per token step (dram__bytes_read.sum + dram__bytes_write.sum of the kernel / tokens)',
        'traffic_source': traffic_src,
        'algorithmic_bytes_per_step': step_bytes, 'flop_per_token': FLOP_PER_TOKEN,
        'phases': [p_ for p_ in phases if p_],
        'share_of_step': {k: v[0] / ms_dev for k, v in prof.items()},
        'phase_us_per_token': us,
        'hbm_bound_tokens_per_s_per_gpu': hbm_tok_s, 'tensor_bound_tokens_per_s_per_gpu': tensor_tok_s,
    }
    line = {
        'metric': METRIC, 'value': value, 'unit': 'tokens/s', 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': ms_dev / args.steps, 'higher_is_better': True, 'scaling': 'strong',
        'vs_baseline': None, 'dtype': 'bf16' if args.precision == 'bf16' else 'f32', 'data': 'synthetic',
        'config': config, 'clocks': clk,
        'e2e': {'value': e2e, 'unit': 'tokens/s', 'h2d_bytes_per_step': prompt_host.numel() * 8 * world,
                'd2h_bytes_per_step': out_host.numel() * 8 * world, 'ms_per_step': ms_e2e / args.steps},
        'gpu_launches': launches, 'roofline': roofline,
    }
    if world > 1 and not args.no_latency:
        # the weak point beside the strong split: every GPU runs the full 1024 streams (what round 1 reported as the headline)
        Bw = args.streams
        pw = torch.randint(6, V, (PROMPT, Bw), generator=g).to(dev)
        ow = torch.empty(T, Bw, dtype=torch.int64, device=dev)
        for i in range(2):
            eng.generate(pw, T, sp(700 + i), out=ow, sync=False)
        ms_w = 0.0
        barrier()
        for i in range(3):
            flush_l2(flush)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            eng.generate(pw, T, sp(710 + i), out=ow, sync=False)
            e1.record(stream)
            torch.cuda.synchronize(dev)
            ms_w += e0.elapsed_time(e1)
        barrier()
        t = torch.tensor([ms_w], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        line['weak_point'] = {'streams_per_gpu': Bw, 'tokens_per_s_whole_box': 3 * Bw * T * world / (float(t.item()) / 1e3),
                              'ms_per_step': float(t.item()) / 3, 'note': 'weak scaling: 1024 streams on every GPU, max over ranks'}
    if rank == 0 and world == 1 and not args.no_latency:
        # p50 latency at batch 1 (BASELINE.json configs[0]: 20-token prompt, 100 new tokens)
        p1 = torch.randint(6, V, (20, 1), generator=g).pin_memory()
        o1 = torch.empty(100, 1, dtype=torch.int64).pin_memory()
        lat = []
        for i in range(103):                 # 3 warm-up + 100 requests
            t0 = time.perf_counter()
            eng.generate(p1, 100, sp(500 + i), out=o1)
            lat.append((time.perf_counter() - t0) * 1e3)
        lat = sorted(lat[3:])
        line['latency_b1'] = {'p50_ms': lat[len(lat) // 2], 'p90_ms': lat[(len(lat) * 9) // 10], 'requests': len(lat), 'workload': 'C1: 20-token prompt + 100 new tokens, batch 1, host in/out',
                              'tokens_per_s': 100 / (lat[len(lat) // 2] / 1e3)}
        line['latency_b1_route'] = route_latency(dev, local_rank, args.precision)
        # BASELINE.json configs[1]: prompt scoring, bptt 70 x batch 256, forward only (logits materialised)
        x2 = torch.randint(0, V, (70, 256), generator=g).to(dev)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(3):
            if it == 1:
                ev0.record(stream)
            eng.reset(256)
            z2, _ = eng.forward(x2, need_raw=False)
            del z2
        ev1.record(stream)
        torch.cuda.synchronize(dev)
        c2_ms = ev0.elapsed_time(ev1) / 2
        line['forward_c2'] = {'ms': c2_ms, 'tokens_per_s': 70 * 256 / (c2_ms / 1e3),
                              'workload': 'C2: forward only, bptt 70 x batch 256, 4.3 GB of fp32 logits written'}
    if rank == 0 and world == 1 and not args.no_latency and B == STREAMS:
        # BASELINE.json configs[2] is 1024 streams IN TOTAL split over 1/2/4/8 GPUs; `--gpus N` under torchrun measures
        # that split on N GPUs.  These are the same per-GPU operating points measured on this one GPU (no collective:
        # the whole box is N times this), L2 flushed between runs as for the headline, each against the HBM roofline of
        # bytes(B) = 88.5 MB + B * 44.0 KB per step.
        pts = []
        for n_gpu in (2, 4, 8):
            b = STREAMS // n_gpu
            pd = prompt_dev[:, :b].contiguous()
            od = torch.empty(T, b, dtype=torch.int64, device=dev)
            for i in range(2):
                eng.generate(pd, T, sp(300 + i), out=od, sync=False)
            torch.cuda.synchronize(dev)
            ms_b = 0.0
            for i in range(3):
                flush_l2(flush)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                eng.generate(pd, T, sp(310 + i), out=od, sync=False)
                e1.record(stream)
                torch.cuda.synchronize(dev)
                ms_b += e0.elapsed_time(e1)
            tok_s = 3 * b * T / (ms_b / 1e3)
            hbm_tok_s = b / ((2 * 44271600 + b * (43200 + 2 * 400 + 8)) / (hbm * 1e9))
            pts.append({'n_gpus': n_gpu, 'streams_per_gpu': b, 'tokens_per_s_per_gpu': tok_s, 'whole_box_tokens_per_s': tok_s * n_gpu,
                        'hbm_roofline_tokens_per_s_per_gpu': hbm_tok_s, 'frac_of_hbm_roofline': tok_s / hbm_tok_s})
        line['strong_split_points'] = pts
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        val, ms, cores = run_cpu_port(1, 1, 4, args.streams)
        line['cpu_baseline'] = {'value': val, 'unit': 'tokens/s', 'cores': cores, 'kind': 'port',
                                'sample': f'{args.streams} streams x 4 new tokens after an untimed prompt (1 warm-up + 1 timed run), torch CPU fp32'}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == '__main__':
    sys.exit(main())
