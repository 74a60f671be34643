import torch
x = torch.empty(1075200000, dtype=torch.float32, device='cuda')   # 4.3 GB
for _ in range(2): x.fill_(1.0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): x.fill_(2.0)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f'fill 4.3 GB: {ms*1e3:.0f} us -> {x.numel()*4/ms/1e6:.0f} GB/s write')
y = torch.empty_like(x[:x.numel()//2]); 
e0.record()
for _ in range(5): y.copy_(x[:x.numel()//2])
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f'copy 2.15 GB: {ms*1e3:.0f} us -> {2*y.numel()*4/ms/1e6:.0f} GB/s read+write')
