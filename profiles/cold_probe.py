"""Does it matter whether the weights of a single-wave GEMM (one LSTM phase) come from DRAM or from L2?
Times awd_tc_gemm_bf16 back to back (weights L2-warm) and with the L2
flushed before every launch (weights cold), CUDA events around each launch."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'pytorch-zappa-serverless_b200'))
import torch
from awd_b200 import _lib
lib = _lib.load()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for M, N, K in [(1024, 4736, 2304), (1024, 4736, 1600), (128, 4736, 2304), (1024, 60000, 448)]:
    A = torch.randn(M, K, device='cuda').to(torch.bfloat16)
    W = torch.randn(N, K, device='cuda').to(torch.bfloat16)
    Cout = torch.empty(M, N, device='cuda')
    def run():
        _lib.check(lib.awd_tc_gemm_bf16(C.c_void_p(A.data_ptr()), K, C.c_void_p(W.data_ptr()), K, M, N, K, None,
                                        C.c_void_p(Cout.data_ptr()), N, st))
    for _ in range(3): run()
    res = {}
    for mode in ('warm', 'cold'):
        ts = []
        for _ in range(20):
            if mode == 'cold':
                flush.add_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); run(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort(); res[mode] = ts[len(ts) // 2]
    print(f'M={M} N={N} K={K}: warm {res["warm"]:.1f} us, cold {res["cold"]:.1f} us per launch', flush=True)
