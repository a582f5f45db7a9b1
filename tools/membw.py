import torch
def t(fn,n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s,e=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e)/n
for mb in (256, 1024):
    x=torch.randn(mb*1024*1024//2,device='cuda').bfloat16(); y=torch.empty_like(x)
    by=x.numel()*2
    a=t(lambda: y.copy_(x)); b=t(lambda: x.mul_(1.5)); c=t(lambda: x.zero_()); d=t(lambda: x.sum())
    print(f"{mb}MB copy {2*by/a/1e6:.0f} GB/s  inplace mul {2*by/b/1e6:.0f} GB/s  memset {by/c/1e6:.0f} GB/s  read(sum) {by/d/1e6:.0f} GB/s")
