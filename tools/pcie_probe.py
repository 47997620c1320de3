import torch, time
x=torch.empty(78643200,dtype=torch.uint8).pin_memory(); d=torch.empty_like(x,device='cuda')
y=torch.empty(14795776,dtype=torch.uint8).pin_memory(); dy=torch.empty_like(y,device='cuda')
for n in (78643200, 9830400, 2457600):
    for _ in range(3): d[:n].copy_(x[:n],non_blocking=True)
    torch.cuda.synchronize(); t=time.perf_counter()
    for _ in range(20): d[:n].copy_(x[:n],non_blocking=True)
    torch.cuda.synchronize(); dt=(time.perf_counter()-t)/20
    print('H2D',n,'bytes %.1f GB/s'%(n/dt/1e9))
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(20): y.copy_(dy,non_blocking=True)
torch.cuda.synchronize(); dt=(time.perf_counter()-t)/20
print('D2H %.1f GB/s'%(14795776/dt/1e9))
