import sys, time, torch
sys.path.insert(0, "/root/repo")
from lgattn_pkg import pkg
B,H,N,M,D = 1,4,1024,1024,64
CALLS = 128
qh = torch.randn(CALLS,B,H,N,D).half().pin_memory(); kh = torch.randn(CALLS,B,H,M,D).half().pin_memory(); vh = torch.randn(CALLS,B,H,M,D).half().pin_memory()
oh = torch.empty(CALLS,B,H,N,D,dtype=torch.float16).pin_memory()
sess = pkg.HostSession(B,H,N,M,pkg.F16_F16,n_slots=4,device=0)
qs = qh[0].numel()*2
def step():
    t0 = time.perf_counter()
    for i in range(CALLS):
        sess.submit(qh.data_ptr()+i*qs, kh.data_ptr()+i*qs, vh.data_ptr()+i*qs, oh.data_ptr()+i*qs, B, N, M)
    t1 = time.perf_counter()
    sess.sync()
    t2 = time.perf_counter()
    return (t1-t0)*1e6/CALLS, (t2-t0)*1e6/CALLS
for _ in range(3): step()
print("per problem: host issue us, total us:", [tuple(round(x,1) for x in step()) for _ in range(4)])
# raw copy bandwidth for 512 KiB and 1.5 MiB pinned copies
d = torch.empty(CALLS*3*qs//2, dtype=torch.float16, device="cuda")
big = torch.empty(CALLS*3*qs//2, dtype=torch.float16).pin_memory()
for chunk in (qs, 3*qs, 12*qs):
    n = (CALLS*3*qs)//chunk
    torch.cuda.synchronize(); t0=time.perf_counter()
    for i in range(n):
        d[i*chunk//2:(i+1)*chunk//2].copy_(big[i*chunk//2:(i+1)*chunk//2], non_blocking=True)
    torch.cuda.synchronize(); dt=time.perf_counter()-t0
    print(f"H2D chunk {chunk/2**20:.2f} MiB: {CALLS*3*qs/dt/1e9:.1f} GB/s")
