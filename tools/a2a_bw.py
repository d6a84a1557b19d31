import os, torch, torch.distributed as dist, time
rank=int(os.environ["RANK"]); world=int(os.environ["WORLD_SIZE"]); torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
for mb in (16, 109, 218):
    n = mb * 1024 * 1024 // 2
    x = torch.randn(n, device="cuda").bfloat16(); y = torch.empty_like(x)
    for _ in range(3): dist.all_to_all_single(y, x)
    torch.cuda.synchronize(); dist.barrier()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): dist.all_to_all_single(y, x)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)/10
    if rank==0: print(f"a2a total {mb} MB per rank (sends {(world-1)/world*mb:.0f} MB out): {ms:.3f} ms -> {(world-1)/world*mb/1024/ms*1000:.1f} GB/s out per GPU")
# peer copy
if rank == 0:
    a = torch.empty(256*1024*1024, dtype=torch.uint8, device="cuda:0")
    try:
        b = torch.empty(256*1024*1024, dtype=torch.uint8, device="cuda:1")
        for _ in range(2): b.copy_(a)
        torch.cuda.synchronize()
        t=time.time(); 
        for _ in range(5): b.copy_(a)
        torch.cuda.synchronize(); dt=(time.time()-t)/5
        print("peer copy 256MB:", round(0.25/dt,1), "GB/s", "can_access_peer", torch.cuda.can_device_access_peer(0,1))
    except Exception as e: print("peer copy failed", e)
dist.destroy_process_group()
