import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloudtik_b200 import ops
from oracle import ops_numpy as ref
for B, F, D in ((2, 27, 128), (1000, 27, 128), (333, 27, 64), (4, 3, 32)):
    rs = np.random.RandomState(B)
    Tn = rs.randn(B, F, D).astype(np.float32)
    T64 = Tn.astype(np.float64)
    want = ref.interact_fwd(T64[:, 0], [T64[:, k] for k in range(1, F)])
    R = ops.interact_fwd(torch.tensor(Tn).cuda()).cpu().numpy().astype(np.float64)
    err = np.abs(R - want)
    print(B, F, D, "max abs err", err.max(), "max |want|", np.abs(want).max(), "mean abs err", err.mean(),
          "bad(>1e-4)", int((err > 1e-4).sum()), "of", err.size)
    if err.max() > 1e-3:
        b = np.unravel_index(err.argmax(), err.shape); print(" worst at", b, R[b], want[b])
