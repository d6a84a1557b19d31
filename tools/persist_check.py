"""Compare the sharded forward (W emulated destination buffers on one GPU) with the plain forward at bench-like sizes."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloudtik_b200 import ops
W, lN, D, P = 4, 8192, 128, int(os.environ.get("P", "20"))
B = W * lN
rows = [39043, 17289, 7420]
T, F, feat0 = len(rows), 5, 1
g = torch.Generator(device="cuda"); g.manual_seed(1)
tw = [torch.rand(n, D, device="cuda", generator=g) - 0.5 for n in rows]
ts = ops.TableSet(tw)
idx = torch.cat([torch.randint(0, n, (B * P,), device="cuda", generator=g) for n in rows])
off = torch.arange(T * B + 1, device="cuda", dtype=torch.int64) * P
for dtype in (torch.bfloat16, torch.float32):
    plain = torch.empty(B, T, D, dtype=dtype, device="cuda")
    ops.emb_fwd(ts, idx, off, B, plain, D, T * D)
    for acc in (False, True):
        for fd in (0, 1, 3):
            bufs = [torch.zeros(lN, F, D, dtype=dtype, device="cuda") for _ in range(W)]
            ptrs = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device="cuda")
            ops.emb_fwd_sharded(ts, idx, off, B, ptrs.data_ptr(), lN, F, feat0, dtype, accumulate=acc, first_dest=fd)
            torch.cuda.synchronize()
            got = torch.cat([b[:, feat0:feat0 + T] for b in bufs], 0)
            err = (got.float() - plain.float()).abs().max().item()
            print(f"persist={os.environ.get('B200_FWD_PERSIST')} dtype={dtype} accumulate={acc} first_dest={fd}: max |diff| {err:.3e}  (|plain| max {plain.float().abs().max().item():.2f})")
