import torch
x = torch.empty(1<<30, dtype=torch.uint8, device='cuda')
y = torch.empty(1<<30, dtype=torch.uint8, device='cuda')
def t(fn, n=10):
    fn(); torch.cuda.synchronize()
    best=1e9
    for _ in range(n):
        a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); best=min(best,a.elapsed_time(b))
    return best
ms=t(lambda: x.zero_()); print("fill 1GiB: %.3f ms -> %.0f GB/s write"%(ms, (1<<30)/ms/1e6))
ms=t(lambda: y.copy_(x)); print("copy 1GiB: %.3f ms -> %.0f GB/s r+w"%(ms, 2*(1<<30)/ms/1e6))
xs = x[:123*1000*1000]
ms=t(lambda: xs.zero_()); print("fill 123MB: %.3f ms -> %.0f GB/s write"%(ms, 123e6/ms/1e6))
s = x.view(torch.float32)
ms=t(lambda: s.sum()); print("read 1GiB: %.3f ms -> %.0f GB/s read"%(ms, (1<<30)/ms/1e6))
