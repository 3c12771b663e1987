import torch, time
N=6400000
h_in=torch.empty(N*3,dtype=torch.float32).pin_memory()
d_in=torch.empty(N*3,dtype=torch.float32,device='cuda')
h_out=torch.empty(N*15,dtype=torch.float64).pin_memory()
d_out=torch.empty(N*15,dtype=torch.float64,device='cuda')
s1=torch.cuda.Stream(); s2=torch.cuda.Stream()
def t(fn,n=5):
    fn(); torch.cuda.synchronize()
    t0=time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return 1e3*(time.perf_counter()-t0)/n
def h2d():
    with torch.cuda.stream(s1): d_in.copy_(h_in,non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h_out.copy_(d_out,non_blocking=True)
def both():
    h2d(); d2h()
print("h2d 77MB %.2f ms"%t(h2d)); print("d2h 768MB %.2f ms"%t(d2h)); print("both concurrently %.2f ms"%t(both))
# bigger concurrent: 768MB both ways
h_in2=torch.empty(N*15,dtype=torch.float64).pin_memory(); d_in2=torch.empty(N*15,dtype=torch.float64,device='cuda')
def h2d2():
    with torch.cuda.stream(s1): d_in2.copy_(h_in2,non_blocking=True)
def both2():
    h2d2(); d2h()
print("h2d 768MB %.2f ms"%t(h2d2)); print("768MB each way concurrently %.2f ms"%t(both2))
