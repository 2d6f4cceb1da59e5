import torch
x=torch.empty(2*1024**3,dtype=torch.uint8,device='cuda'); y=torch.empty_like(x)
def t(fn,n=5):
    fn(); ts=[]
    for _ in range(n):
        a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts)
print('fill 2GB  GB/s', 2*1.0737/ (t(lambda: x.zero_())*1e-3))
print('copy 2GB r+w GB/s', 4*1.0737/(t(lambda: y.copy_(x))*1e-3))
print('read-sum GB/s', 2*1.0737/(t(lambda: x.view(torch.int32).sum())*1e-3))
