import torch
dev=torch.device('cuda',0)
flush=torch.empty(256*1024*1024,dtype=torch.uint8,device=dev)
def t(fn,n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ts=[]
    for _ in range(n):
        flush.fill_(1)
        e0=torch.cuda.Event(enable_timing=True);e1=torch.cuda.Event(enable_timing=True)
        e0.record();fn();e1.record();e1.synchronize();ts.append(e0.elapsed_time(e1))
    ts.sort();return ts[len(ts)//2],ts[0]
for mb in (768,1536,4096):
    x=torch.empty(mb*1024*1024//8,dtype=torch.float64,device=dev)
    m,mn=t(lambda:x.fill_(1.5))
    print('fill %d MiB: median %.4f ms min %.4f -> %.0f GB/s'%(mb,m,mn,mb*1.048576/m))
    m,mn=t(lambda:torch.cuda.current_stream().synchronize() or x.zero_())
    print('zero %d MiB: median %.4f ms -> %.0f GB/s'%(mb,m,mb*1.048576/m))
a=torch.empty(768*1024*1024//8,dtype=torch.float64,device=dev);b=torch.empty_like(a)
m,mn=t(lambda:b.copy_(a))
print('copy 768 MiB (r+w 1536): median %.4f -> %.0f GB/s'%(m,2*768*1.048576/m))
s=torch.empty(768*1024*1024//8//8,dtype=torch.float64,device=dev)
m,mn=t(lambda:s.sum())
r=torch.empty(768*1024*1024//8,dtype=torch.float64,device=dev)
m,mn=t(lambda:r.sum())
print('read 768 MiB: median %.4f -> %.0f GB/s'%(m,768*1.048576/m))
