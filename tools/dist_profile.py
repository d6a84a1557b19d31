"""torch.profiler kernel table of the distributed bench step on rank 0 (eager, a few steps).
    python -m torch.distributed.run --nproc-per-node 2 ... tools/dist_profile.py"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from cloudtik_b200 import ext_dist
from cloudtik_b200.dlrm import DLRM_Net

world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"]); lr_ = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr_); dev = torch.device("cuda", lr_)
ext_dist.init_distributed(rank=rank, size=world, backend="nccl")
rows = bench.table_rows(0)
np.random.seed(727)
net = DLRM_Net(bench.D, np.array(rows), np.array(bench.LN_BOT), np.array(bench.LN_TOP), sigmoid_top=len(bench.LN_TOP) - 2,
               device=dev, lr=1.0 / world, bf16=True)
dense = [p for p in net.parameters() if p.requires_grad]
opt = torch.optim.SGD(dense, lr=1.0)
sync = ext_dist.DenseGradAllreduce(dense)
a = type("A", (), {"sets": 2})()
B = 32768
sets = bench.make_batches(a, [rows[i] for i in net.local_ln_emb], B, B * world, dev, rank)
def step(i):
    X, Y, idx, off = sets[i % 2]["dev"]
    E = net.loss_fn(net(X, idx, off, None).float(), Y)
    opt.zero_grad(set_to_none=True); E.backward(); sync(); opt.step()
for i in range(5): step(i)
torch.cuda.synchronize(); torch.distributed.barrier()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(4): step(i)
    torch.cuda.synchronize()
if rank == 0:
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=28, max_name_column_width=70))
torch.distributed.destroy_process_group()
