import torch, time
dev=torch.device("cuda",0)
for mb in (3, 12, 96):
    n=mb*1024*1024
    h=torch.empty(n,dtype=torch.uint8).pin_memory(); d=torch.empty(n,dtype=torch.uint8,device=dev)
    h2=torch.empty(n//2,dtype=torch.uint8).pin_memory(); d2=torch.empty(n//2,dtype=torch.uint8,device=dev)
    s1=torch.cuda.Stream(); s2=torch.cuda.Stream()
    for _ in range(3): d.copy_(h,non_blocking=True)
    torch.cuda.synchronize()
    reps=max(4,int(2000/mb)); t=time.perf_counter()
    for _ in range(reps): d.copy_(h,non_blocking=True)
    torch.cuda.synchronize(); dt=time.perf_counter()-t
    a=n*reps/dt/1e9
    t=time.perf_counter()
    for _ in range(reps):
        with torch.cuda.stream(s1): d.copy_(h,non_blocking=True)
        with torch.cuda.stream(s2): h2.copy_(d2,non_blocking=True)
    torch.cuda.synchronize(); dt=time.perf_counter()-t
    print(f"{mb} MB: H2D alone {a:.1f} GB/s; with concurrent D2H(half size): H2D {n*reps/dt/1e9:.1f} GB/s + D2H {n/2*reps/dt/1e9:.1f} GB/s")
