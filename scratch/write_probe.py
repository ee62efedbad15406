import torch, json
def timed(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b)/it
n=2048*224*224*3
x=torch.randint(0,255,(n,),dtype=torch.uint8,device='cuda'); y=torch.empty(n,dtype=torch.bfloat16,device='cuda')
t=timed(lambda: y.copy_(x)); print('u8->bf16 convert', round(t,4),'ms', round(n*3/t/1e6),'GB/s')
t=timed(lambda: y.zero_()); print('bf16 fill', round(t,4),'ms', round(n*2/t/1e6),'GB/s')
z=torch.empty_like(y); t=timed(lambda: z.copy_(y)); print('bf16 copy', round(t,4),'ms', round(n*4/t/1e6),'GB/s')
