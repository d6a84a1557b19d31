"""Error of the fp32 interaction forward (set B200_INTERACT_TF32=1 for the 3xTF32 tcgen05 kernel) against float64."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloudtik_b200 import ops
for B, F, D in ((2, 27, 128), (1000, 27, 128), (333, 27, 64), (4, 3, 32)):
    rs = np.random.RandomState(B)
    Tn = rs.randn(B, F, D).astype(np.float32)
    T64 = Tn.astype(np.float64)
    Z = np.matmul(T64, T64.transpose(0, 2, 1))  # float64 reference of the pairwise dots, written out here:
    li, lj = np.array([(i, j) for i in range(F) for j in range(i)]).T  # a tool must not reach into oracle/
    want = np.concatenate([T64[:, 0], Z[:, li, lj]], axis=1)
    R = ops.interact_fwd(torch.tensor(Tn).cuda()).cpu().numpy().astype(np.float64)
    err = np.abs(R - want)
    print(B, F, D, "max abs err", err.max(), "max |want|", np.abs(want).max(), "mean abs err", err.mean(),
          "bad(>1e-4)", int((err > 1e-4).sum()), "of", err.size)
    if err.max() > 1e-3:
        b = np.unravel_index(err.argmax(), err.shape); print(" worst at", b, R[b], want[b])
