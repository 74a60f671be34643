#!/usr/bin/env python
"""Micro-benchmark of awd_tc_gemm_bf16 (tcgen05 GEMM, store epilogue) on the engine's shapes.
Repeats each launch back to back so the weights stay L2-resident; prints us/launch and TFLOP/s.
Knob via env: AWD_TC_PAIR."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))
import torch
from awd_b200 import _lib
lib = _lib.load()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
shapes = [(1024, 4600, 1600), (1024, 4600, 2304), (1024, 1600, 1600), (1024, 15360, 448), (128, 4600, 1600), (17920, 4600, 448)]
if os.environ.get('MICRO_SWEEP'):
    shapes = [(M, 4600, K) for M in (128, 1024) for K in (64, 448, 1600, 3200, 6400)] + [(1024, 4736, 1600), (1024, 4736, 6400)]
if os.environ.get('MICRO_SHAPES'):
    shapes = [tuple(int(v) for v in t.split('x')) for t in os.environ['MICRO_SHAPES'].split(',')]
for M, N, K in shapes:
    A = torch.randn(M, K, device='cuda').to(torch.bfloat16)
    W = torch.randn(N, K, device='cuda').to(torch.bfloat16)
    Cout = torch.empty(M, N, device='cuda')
    def run():
        _lib.check(lib.awd_tc_gemm_bf16(C.c_void_p(A.data_ptr()), K, C.c_void_p(W.data_ptr()), K, M, N, K, None,
                                        C.c_void_p(Cout.data_ptr()), N, st))
    for _ in range(5): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 50
    e0.record()
    for _ in range(n): run()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    print(f'M={M:6d} N={N:6d} K={K:5d}  {us:8.1f} us/launch  {2*M*N*K/us/1e6:8.1f} TFLOP/s', flush=True)
