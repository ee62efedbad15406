import os, sys, time, torch, torch.distributed as dist
rank=int(os.environ["RANK"]); local=int(os.environ["LOCAL_RANK"]); world=int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
if rank==0: print("peer access 0->1:", torch.cuda.can_device_access_peer(0,1), flush=True)
for nbytes in (8, 1<<16, 3276800, 1<<25):
    x=torch.empty(nbytes//8, dtype=torch.int64, device="cuda"); out=torch.empty(world*(nbytes//8), dtype=torch.int64, device="cuda")
    for _ in range(5): dist.all_gather_into_tensor(out, x)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): dist.all_gather_into_tensor(out, x)
    e1.record(); torch.cuda.synchronize()
    if rank==0: print(f"all_gather {nbytes} B/rank: {e0.elapsed_time(e1)/20*1e3:.1f} us", flush=True)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1024, dtype=torch.int64, device="cuda")
    h = symm.rendezvous(t, dist.group.WORLD)
    if rank==0: print("symm_mem ok:", type(h), [m for m in dir(h) if not m.startswith('_')][:40], flush=True)
    peer = h.get_buffer((rank+1)%world, (1024,), torch.int64)
    t.fill_(rank+1); h.barrier()
    print(rank, "peer value", int(peer[0].item()), flush=True)
    h.barrier()
except Exception as e:
    print("symm_mem failed:", repr(e), flush=True)
dist.destroy_process_group()
