import torch, time
n=512*307200
h=torch.empty(n,dtype=torch.uint8,pin_memory=True); d=torch.empty(n,dtype=torch.uint8,device='cuda')
ho=torch.empty(32*1024*1024,dtype=torch.uint8,pin_memory=True); do=torch.empty(32*1024*1024,dtype=torch.uint8,device='cuda')
s1=torch.cuda.Stream(); s2=torch.cuda.Stream()
for chunks in (1,16):
    for both in (False,True):
        torch.cuda.synchronize(); 
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        best=1e9
        for rep in range(5):
            torch.cuda.synchronize(); t=time.perf_counter()
            with torch.cuda.stream(s1):
                c=n//chunks
                for i in range(chunks): d[i*c:(i+1)*c].copy_(h[i*c:(i+1)*c],non_blocking=True)
            if both:
                with torch.cuda.stream(s2): ho.copy_(do,non_blocking=True)
            torch.cuda.synchronize(); best=min(best,time.perf_counter()-t)
        print("chunks",chunks,"with D2H" if both else "H2D only","%.3f ms  %.1f GB/s"%(best*1e3,n/best/1e9))
