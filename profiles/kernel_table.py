#!/usr/bin/env python
"""Per-kernel throughput table on the shipped sizes (CUDA events, back-to-back launches, inputs > L2
where the kernel streams).  Used for profiles/r01_kernel_table.md; also the command ncu wraps."""
import ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))
import numpy as np, torch
from awd_b200 import Engine, SamplingParams, _lib
from bench import weights_numpy, V, E, H, L

def timed(fn, n=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e-3

rows = []
sd = weights_numpy(0)
quick = bool(os.environ.get('KT_QUICK'))
for prec in ('bf16', 'fp32'):
    eng = Engine(V, E, H, L, precision=prec)
    eng.load_state_dict(sd)
    # K1 gather: 70*256*16 tokens, output 459 MB fp32 (> L2)
    tok = torch.randint(0, V, (70 * 256 * 16,), device='cuda')
    t = timed(lambda: eng.gather(tok), n=3 if quick else 10)
    bytes_ = tok.numel() * (E * (2 if prec == 'bf16' else 4) + E * 4 + 8)
    rows.append((f'K1 gather ({prec} table -> fp32 rows, {tok.numel()} tokens)', t * 1e6, f'{bytes_ / t / 1e9:.0f} GB/s'))
    # forward C2: bptt 70 x batch 256, logits materialised (4.3 GB)
    x = torch.randint(0, V, (70, 256), device='cuda')
    eng.reset(256)
    t = timed(lambda: eng.forward(x, need_raw=False), n=2 if quick else 5, warm=1)
    rows.append((f'forward C2 (70x256, logits 4.3 GB) {prec}', t * 1e6, f'{70 * 256 / t:.0f} tokens/s, {70 * 256 * 88.38e6 / t / 1e12:.1f} TFLOP/s'))
    t = timed(lambda: eng.forward(x, need_raw=False, last_step_only=True), n=2 if quick else 5, warm=1)
    rows.append((f'forward C2, last-step logits only {prec}', t * 1e6, f'{70 * 256 / t:.0f} tokens/s'))
    # sampler on materialised logits (fp32 kernel in both modes)
    if prec == 'fp32':
        z = torch.randn(1024, V, device='cuda')
        sp = SamplingParams(banned=[3, 4, 5], seed=1)
        t = timed(lambda: eng.sample(z, sp, 0))
        rows.append(('sample_f32_kernel (1024 x 60000 logits, Philox race)', t * 1e6, f'{z.numel() * 4 / t / 1e9:.0f} GB/s'))
    if prec == 'bf16':
        p1 = torch.randint(6, V, (20, 1)).pin_memory(); o1 = torch.empty(100, 1, dtype=torch.int64).pin_memory()
        sp = SamplingParams(banned=[3, 4, 5], seed=1)
        t = timed(lambda: eng.generate(p1, 100, sp, out=o1), n=3 if quick else 10)
        rows.append(('lat_generate_kernel C1 (20-token prompt + 100 tokens, batch 1, host in/out)', t * 1e6,
                     f'{t * 1e6 / 120:.1f} us/token; weights streamed per token 66 MB -> {66e6 * 120 / t / 1e9:.0f} GB/s'))
    eng.close()
lib = _lib.load()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for (M, N, K, name) in [(17920, 4600, 448, 'K2 input projection, layer 0, whole C2 prompt'), (17920, 4600, 1152, 'K2 input projection, layer 1'),
                        (17920, 60000, 448, 'K4 decoder logits for forward() (4.3 GB out)')]:
    A = torch.randn(M, K, device='cuda').to(torch.bfloat16); W = torch.randn(N, K, device='cuda').to(torch.bfloat16)
    Cout = torch.empty(M, N, device='cuda')
    f = lambda: _lib.check(lib.awd_tc_gemm_bf16(C.c_void_p(A.data_ptr()), K, C.c_void_p(W.data_ptr()), K, M, N, K, None, C.c_void_p(Cout.data_ptr()), N, st))
    t = timed(f, n=3 if quick else 10)
    rows.append((f'tc_gemm_kernel {name} ({M}x{N}x{K})', t * 1e6, f'{2 * M * N * K / t / 1e12:.0f} TFLOP/s, C write {M * N * 4 / t / 1e9:.0f} GB/s'))
    del A, W, Cout
for r in rows:
    print(f'| {r[0]} | {r[1]:.1f} | {r[2]} |', flush=True)
