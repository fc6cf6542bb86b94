import torch,time
a=torch.empty(1<<30,dtype=torch.bfloat16,device='cuda'); b=torch.empty_like(a)
for _ in range(3): b.copy_(a)
torch.cuda.synchronize()
best=1e9
for _ in range(10):
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize()
    best=min(best,e0.elapsed_time(e1))
print("copy GB/s", 2*a.numel()*2/best/1e6)
