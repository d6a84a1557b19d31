import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank=int(os.environ["RANK"]); world=int(os.environ["WORLD_SIZE"]); torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
t = symm.empty(1024, dtype=torch.float32, device="cuda")
hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "sig", [hex(p) for p in hdl.signal_pad_ptrs][:2], "ws", hdl.world_size, flush=True)
t.fill_(rank + 1.0)
hdl.barrier()
peer = hdl.get_buffer((rank + 1) % world, (1024,), torch.float32)
print(rank, "peer value", float(peer[0].item()), flush=True)
hdl.barrier()
print(rank, "attrs", [a for a in dir(hdl) if not a.startswith("_")], flush=True)
dist.destroy_process_group()
