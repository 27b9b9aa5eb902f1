import torch, time
x=torch.empty(1<<28, dtype=torch.float32, device='cuda')  # 1 GiB
y=torch.empty_like(x)
def t(f,n=10):
    f(); torch.cuda.synchronize()
    s=torch.cuda.Event(enable_timing=True); e=torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): f()
    e.record(); torch.cuda.synchronize(); return s.elapsed_time(e)/n
b=x.numel()*4
ms=t(lambda: x.fill_(1.0)); print('fill  GB/s', b/ms/1e6)
ms=t(lambda: y.copy_(x)); print('copy  GB/s (r+w)', 2*b/ms/1e6)
ms=t(lambda: x.sum()); print('read  GB/s', b/ms/1e6)
ms=t(lambda: torch.add(x, 1.0, out=y)); print('add   GB/s (r+w)', 2*b/ms/1e6)
