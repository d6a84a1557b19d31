"""Generate the known-answer fixtures under ``tests/golden/`` FROM THE REFERENCE ITSELF.

TEST INFRASTRUCTURE.  Runs only in the build container (needs ``/root/reference``);
the fixtures it writes are committed so the GPU box never needs the reference.

    python -m oracle.gen_golden            # rewrites tests/golden/*.npz

What is imported from the reference (read-only, unmodified):
  * ``P/dlrm_s_pytorch.py``  -> ``DLRM_Net`` (model, interaction), ``LRPolicyScheduler``
  * ``P/extend_distributed.py`` -> ``get_my_slice`` / ``get_split_lengths``
  * ``T/dlrm_data_pytorch.py`` -> ``generate_uniform_input_batch`` (ragged multi-hot bags)
  * ``T/data_loader_terabyte.py`` -> ``_transform_features`` (binary Criteo records -> model inputs)
  * ``T/tricks/{qr,md}_embedding_bag.py`` -> ``QREmbeddingBag``, ``PrEmbeddingBag``, ``md_solver``
  * ``T/dlrm_s_pytorch.py`` + ``T/dlrm_data_pytorch.py`` -> the two configurations of the reference's own test
    script ``T/test/dlrm_s_test.sh`` (model, random data loader, MSE / SGD loop)
with ``P/`` = applications/ai/quickstart/models/recommendation/pytorch/dlrm/product and
``T/`` = .../dlrm/training/bfloat16.  Absent third-party modules (mlperf_logging) are
stubbed in ``sys.modules``; nothing in the reference tree is edited.
"""
from __future__ import annotations

import importlib
import importlib.machinery
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/applications/ai/quickstart/models/recommendation/pytorch/dlrm"
P_DIR = os.path.join(REF, "product")
T_DIR = os.path.join(REF, "training", "bfloat16")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

_SHARED_NAMES = ["data_utils", "data_loader_terabyte", "extend_distributed", "dlrm_data_pytorch",
                 "mlperf_logger"]


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, loader=None)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def load_T_generator():
    """T/dlrm_data_pytorch.generate_uniform_input_batch, loaded in isolation."""
    sys.dont_write_bytecode = True
    consts = _stub("mlperf_logging.mllog.constants")
    mllog = _stub("mlperf_logging.mllog", constants=consts,
                  get_mllogger=lambda: types.SimpleNamespace())
    _stub("mlperf_logging", mllog=mllog)
    sys.path.insert(0, T_DIR)
    try:
        mod = importlib.import_module("dlrm_data_pytorch")
    finally:
        sys.path.remove(T_DIR)
    for n in _SHARED_NAMES:  # P/ has modules of the same names
        sys.modules.pop(n, None)
    return mod


def load_P():
    sys.dont_write_bytecode = True
    sys.path.insert(0, P_DIR)
    try:
        M = importlib.import_module("dlrm_s_pytorch")
        X = importlib.import_module("extend_distributed")
    finally:
        sys.path.remove(P_DIR)
    # apply_emb / interact_features read a module-global ``args`` (P/dlrm_s_pytorch.py:288,310)
    M.args = types.SimpleNamespace(ipex=False, ipex_interaction=False)
    return M, X


def _state(model):
    return {k: v.detach().numpy().copy() for k, v in model.state_dict().items()}


def gen_interaction(M):
    """Reference interact_features forward + autograd backward."""
    net = M.DLRM_Net.__new__(M.DLRM_Net)  # interact_features uses no attributes
    out = {}
    x = torch.tensor([[1.0, 2.0]])
    ly = [torch.tensor([[3.0, 4.0]]), torch.tensor([[5.0, 6.0]])]
    out["kat_R"] = M.DLRM_Net.interact_features(net, x, ly).numpy()
    rs = np.random.RandomState(11)
    for tag, (B, F, D) in {"a": (5, 4, 8), "b": (3, 27, 16), "c": (2, 27, 128)}.items():
        x = torch.tensor(rs.randn(B, D).astype(np.float32), requires_grad=True)
        ly = [torch.tensor(rs.randn(B, D).astype(np.float32), requires_grad=True) for _ in range(F - 1)]
        R = M.DLRM_Net.interact_features(net, x, ly)
        dR = torch.tensor(rs.randn(*R.shape).astype(np.float32))
        R.backward(dR)
        out[f"{tag}_x"] = x.detach().numpy()
        out[f"{tag}_ly"] = np.stack([t.detach().numpy() for t in ly])
        out[f"{tag}_R"] = R.detach().numpy()
        out[f"{tag}_dR"] = dR.numpy()
        out[f"{tag}_dx"] = x.grad.numpy()
        out[f"{tag}_dly"] = np.stack([t.grad.numpy() for t in ly])
    np.savez_compressed(os.path.join(OUT, "interaction.npz"), **out)


def gen_index_kats(X):
    """Sparse-grad structure / bucketing KAT + table partition."""
    out = {}
    idx = torch.tensor([3, 1, 3, 7, 7, 7, 0, 9])
    off = torch.tensor([0, 2, 2, 5])
    E = torch.nn.EmbeddingBag(10, 4, mode="sum", sparse=True)
    y = E(idx, off)
    g = torch.arange(16, dtype=torch.float32).view(4, 4)
    y.backward(g)
    G = E.weight.grad
    out["idx"], out["off"], out["gout"] = idx.numpy(), off.numpy(), g.numpy()
    out["coo_indices"] = G._indices().numpy()[0]
    out["coo_values"] = G._values().numpy()
    C = G.coalesce()
    out["coalesced_indices"] = C.indices().numpy()[0]
    out["coalesced_values"] = C.values().numpy()
    s, perm = torch.sort(idx, stable=True)
    u, inv, cnt = torch.unique(idx, sorted=True, return_inverse=True, return_counts=True)
    out["sorted"], out["perm"] = s.numpy(), perm.numpy()
    out["unique"], out["inverse"], out["counts"] = u.numpy(), inv.numpy(), cnt.numpy()
    _, o2b, bag_size, _ = torch._embedding_bag(E.weight.detach(), idx, off, False, 1)  # mode=mean fills offset2bag
    out["offset2bag"], out["bag_size"] = o2b.numpy(), bag_size.numpy()
    # table partition from the reference's own functions, with a fake process group
    rows = []
    for world in (1, 2, 3, 4, 8):
        for rank in range(world):
            X.dist = types.SimpleNamespace(get_world_size=lambda w=world: w, get_rank=lambda r=rank: r)
            for n in (26, 8, 5, 16):
                sl = X.get_my_slice(n)
                my_len, splits = X.get_split_lengths(n)
                rows.append([world, rank, n, sl.start, sl.stop, my_len] +
                            (list(splits) + [-1] * (8 - world) if splits else [-1] * 8))
    out["partition"] = np.array(rows, dtype=np.int64)
    np.savez_compressed(os.path.join(OUT, "index_kats.npz"), **out)


def _run_reference_model(M, seed, D, ln_emb, ln_bot, ln_top, batches, lr):
    """Train the reference DLRM_Net with torch.optim.SGD exactly as the loop at
    P/dlrm_s_pytorch.py:1029-1059 does (fp32, no IPEX)."""
    np.random.seed(seed)
    torch.manual_seed(seed)
    net = M.DLRM_Net(D, np.asarray(ln_emb), np.asarray(ln_bot), np.asarray(ln_top),
                     sigmoid_bot=-1, sigmoid_top=len(ln_top) - 2, np_init_emb_weight=True)
    init = _state(net)
    opt = torch.optim.SGD(net.parameters(), lr=lr)
    losses = []
    for (X, lS_o, lS_i, T) in batches:
        Z = net(X, lS_o, lS_i).float()
        E = net.loss_fn(Z, T)
        losses.append(float(E.detach().cpu().numpy()))
        opt.zero_grad(set_to_none=True)
        E.backward()
        opt.step()
    return net, init, np.array(losses, dtype=np.float64)


def gen_small_model(M, G):
    """Tiny ragged multi-hot model; full init/inputs/final weights stored."""
    seed, D = 123, 8
    ln_emb = [40, 30, 20, 10]
    F = len(ln_emb) + 1
    ln_bot = [13, 16, D]
    ln_top = [D + F * (F - 1) // 2, 16, 1]
    B, steps, lr = 6, 4, 0.1
    G.ra.seed(seed + 1)
    batches = []
    for _ in range(steps):
        X, lS_o, lS_i = G.generate_uniform_input_batch(13, np.asarray(ln_emb), B, 5, False)
        T = G.generate_random_output_batch(B, 1, round_targets=True)
        batches.append((X, lS_o, lS_i, T))
    net, init, losses = _run_reference_model(M, seed, D, ln_emb, ln_bot, ln_top, batches, lr)
    out = {"seed": seed, "D": D, "ln_emb": np.array(ln_emb), "ln_bot": np.array(ln_bot),
           "ln_top": np.array(ln_top), "lr": lr, "losses": losses, "steps": steps, "B": B}
    for k, v in init.items():
        out["init/" + k] = v
    for k, v in _state(net).items():
        out["final/" + k] = v
    for s, (X, lS_o, lS_i, T) in enumerate(batches):
        out[f"X{s}"], out[f"T{s}"] = X.numpy(), T.numpy()
        for t in range(len(ln_emb)):
            out[f"o{s}_{t}"], out[f"i{s}_{t}"] = lS_o[t].numpy(), lS_i[t].numpy()
    np.savez_compressed(os.path.join(OUT, "small_model.npz"), **out)


KAGGLE_ROWS = [1460, 583, 1000000, 1000000, 305, 24, 12517, 633, 3, 93145, 5683, 1000000, 3194, 27,
               14992, 1000000, 10, 5652, 2173, 4, 1000000, 18, 15, 286181, 105, 142572]


def gen_kaggle_c1(M, G):
    """BASELINE config 1 shape (T/bench/dlrm_s_criteo_kaggle.sh:24): 26 tables capped at
    1M rows, D=16, bot 13-512-256-64-16, top 512-256-1, B=128, SGD lr 0.1, BCE, one-hot.
    Weights are regenerated from the seed by the port; inputs + loss trace + a sample of
    the updated rows are stored."""
    seed, D, B, steps, lr = 727, 16, 128, 8, 0.1
    ln_bot = [13, 512, 256, 64, 16]
    F = 27
    ln_top = [D + F * (F - 1) // 2, 512, 256, 1]
    G.ra.seed(seed + 1)
    batches = []
    for _ in range(steps):
        X, lS_o, lS_i = G.generate_uniform_input_batch(13, np.asarray(KAGGLE_ROWS), B, 1, True)
        T = G.generate_random_output_batch(B, 1, round_targets=True)
        batches.append((X, lS_o, lS_i, T))
    net, init, losses = _run_reference_model(M, seed, D, KAGGLE_ROWS, ln_bot, ln_top, batches, lr)
    out = {"seed": seed, "D": D, "ln_emb": np.array(KAGGLE_ROWS), "ln_bot": np.array(ln_bot),
           "ln_top": np.array(ln_top), "lr": lr, "losses": losses, "steps": steps, "B": B}
    out["X"] = np.stack([b[0].numpy() for b in batches])
    out["T"] = np.stack([b[3].numpy() for b in batches])
    out["idx"] = np.stack([np.stack([t.numpy() for t in b[2]]) for b in batches]).astype(np.int32)
    out["off"] = np.stack([np.stack([t.numpy() for t in b[1]]) for b in batches]).astype(np.int32)
    fin = _state(net)
    # pins on the weights: float64 sums of init tables (checks the port's RNG stream) and
    # the rows of table 2 / table 23 touched in the last step after all updates
    out["init_table_sums"] = np.array([init[f"emb_l.{t}.weight"].astype(np.float64).sum()
                                       for t in range(26)])
    out["final_table_sums"] = np.array([fin[f"emb_l.{t}.weight"].astype(np.float64).sum()
                                        for t in range(26)])
    for t in (2, 23):
        rows = np.unique(out["idx"][-1, t])
        out[f"rows_t{t}"] = rows
        out[f"final_rows_t{t}"] = fin[f"emb_l.{t}.weight"][rows]
    out["final_top0_w"] = fin["top_l.0.weight"][:8]
    out["final_bot0_w"] = fin["bot_l.0.weight"]
    np.savez_compressed(os.path.join(OUT, "kaggle_c1.npz"), **out)


def gen_lr_policy(M):
    """LRPolicyScheduler trace (P/dlrm_s_pytorch.py:133-167)."""
    p = torch.nn.Parameter(torch.zeros(1))
    cfgs = [(1.0, 5, 8, 6), (18.0, 2, 3, 4), (0.5, 4, 4, 0)]  # warm-up 0 trips an upstream AttributeError (last_lr unset)
    out = {"cfgs": np.array(cfgs, dtype=np.float64)}
    for c, (base, warm, start, nsteps) in enumerate(cfgs):
        opt = torch.optim.SGD([p], lr=base)
        sch = M.LRPolicyScheduler(opt, warm, start, nsteps)
        trace = []
        for _ in range(20):
            trace.append(sch.get_last_lr()[0])
            opt.step()
            sch.step()
        out[f"trace{c}"] = np.array(trace, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "lr_policy.npz"), **out)


def load_T_terabyte():
    """T/data_loader_terabyte (``_transform_features``), loaded in isolation like the generator."""
    sys.dont_write_bytecode = True
    sys.path.insert(0, T_DIR)
    try:
        mod = importlib.import_module("data_loader_terabyte")
    finally:
        sys.path.remove(T_DIR)
    for n in _SHARED_NAMES:
        sys.modules.pop(n, None)
    return mod


def gen_criteo_transform(TB):
    """Binary-record decode + ``_transform_features`` (T/data_loader_terabyte.py:83-102, :259-271) on
    synthetic int32 records: label, 13 dense counts (incl. 0 and large values), 26 categorical ids
    (incl. negative ones = hashed ids that wrapped, and ids >= max_ind_range)."""
    rs = np.random.RandomState(2024)
    out = {}
    for name, N, mir in (("a", 37, 40_000_000), ("b", 64, 1000), ("c", 5, 0)):
        rec = np.empty((N, 40), dtype=np.int32)
        rec[:, 0] = rs.randint(0, 2, size=N)
        rec[:, 1:14] = rs.randint(0, 3_000_000, size=(N, 13)) * (rs.rand(N, 13) < 0.8)
        rec[:, 14:] = rs.randint(-2**31, 2**31 - 1, size=(N, 26), dtype=np.int64).astype(np.int32)
        if mir == 0:
            rec[:, 14:] = np.abs(rec[:, 14:] // 2)
        t = torch.from_numpy(rec.copy())
        X, lS_o, lS_i, Y = TB._transform_features(x_int_batch=t[:, 1:14], x_cat_batch=t[:, 14:], y_batch=t[:, 0],
                                                  max_ind_range=mir, flag_input_torch_tensor=True)
        out.update({f"{name}_records": rec, f"{name}_max_ind_range": np.int64(mir), f"{name}_X": X.numpy(),
                    f"{name}_lS_o": lS_o.numpy(), f"{name}_lS_i": lS_i.contiguous().numpy(), f"{name}_Y": Y.numpy()})
    np.savez_compressed(os.path.join(OUT, "criteo_transform.npz"), **out)


def gen_int8_tables():
    """Weight half of the reference's int8 QConfig (P/dlrm_s_pytorch.py:431-432,870-871):
    PerChannelMinMaxObserver(dtype=qint8, qscheme=per_channel_symmetric) on an EmbeddingBag weight (channel axis 0
    = one scale per row) + torch.quantize_per_channel -- run here with torch's own implementation."""
    from torch.ao.quantization import PerChannelMinMaxObserver
    rs = np.random.RandomState(77)
    W = (rs.randn(200, 64) * (rs.rand(200, 1) * 2 + 1e-3)).astype(np.float32)
    W[3] = 0.0            # all-zero row: scale = eps
    W[4, :] = 1e-20       # below eps * 127.5
    W[5, 0] = -7.0        # min dominates
    obs = PerChannelMinMaxObserver(ch_axis=0, dtype=torch.qint8, qscheme=torch.per_channel_symmetric)
    obs(torch.tensor(W))
    scale, zp = obs.calculate_qparams()
    q = torch.quantize_per_channel(torch.tensor(W), scale, zp, 0, torch.qint8)
    lens = rs.randint(0, 5, size=50)
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    idx = rs.randint(0, 200, size=int(lens.sum())).astype(np.int64)
    deq = q.dequantize()
    pooled = torch.nn.functional.embedding_bag(torch.tensor(idx), deq, torch.tensor(off[:-1]), mode="sum")
    np.savez_compressed(os.path.join(OUT, "int8_tables.npz"), W=W, scale=scale.float().numpy(),
                        zero_point=zp.numpy(), q=q.int_repr().numpy(), idx=idx, off=off, pooled=pooled.numpy())


def gen_tricks():
    """T/tricks (QREmbeddingBag, PrEmbeddingBag, md_solver) run on the CPU: one forward + backward each, and the
    dimension solver's answers.  Includes ids above 2**24, where the reference's float32 quotient is inexact."""
    sys.dont_write_bytecode = True
    sys.path.insert(0, T_DIR)
    try:
        qr = importlib.import_module("tricks.qr_embedding_bag")
        md = importlib.import_module("tricks.md_embedding_bag")
    finally:
        sys.path.remove(T_DIR)
    rs = np.random.RandomState(31)
    out = {}
    B, D, C = 40, 16, 7
    lens = rs.randint(0, 5, size=B)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    for op in ("mult", "add", "concat"):
        ncat = 40_000_003 if op == "mult" else 1000
        nq = int(np.ceil(ncat / C))
        idx = rs.randint(0, ncat, size=int(lens.sum())).astype(np.int64)
        if op == "mult":
            idx[:4] = [33554433, 16777217, 40_000_002, 25_165_829]  # > 2**24: float32 quotient territory
        # only the rows that are touched get random values (a 5.7 M x 16 table of noise would bloat the fixture)
        wq = np.full((nq, D), 0.25, dtype=np.float32)
        q_rows = np.unique(np.concatenate([(torch.tensor(idx) / C).long().numpy(), idx // C]))
        wq[q_rows] = rs.randn(len(q_rows), D)
        wr = rs.randn(C, D).astype(np.float32)
        m = qr.QREmbeddingBag(ncat, D, C, operation=op, mode="sum", sparse=True,
                              _weight=[torch.tensor(wq), torch.tensor(wr)])
        y = m(torch.tensor(idx), torch.tensor(off))
        gy = rs.randn(*y.shape).astype(np.float32)
        y.backward(torch.tensor(gy))
        gq, gr = m.weight_q.grad.coalesce(), m.weight_r.grad.coalesce()
        out.update({f"qr_{op}_ncat": np.int64(ncat), f"qr_{op}_idx": idx, f"qr_{op}_off": off,
                    f"qr_{op}_q_rows": q_rows, f"qr_{op}_wq_rows": wq[q_rows], f"qr_{op}_wr": wr,
                    f"qr_{op}_y": y.detach().numpy(), f"qr_{op}_gy": gy,
                    f"qr_{op}_gq_idx": gq.indices().numpy()[0], f"qr_{op}_gq_val": gq.values().numpy(),
                    f"qr_{op}_gr_idx": gr.indices().numpy()[0], f"qr_{op}_gr_val": gr.values().numpy()})
    # PrEmbeddingBag
    torch.manual_seed(5)
    pm = md.PrEmbeddingBag(500, 8, 32)
    idx = rs.randint(0, 500, size=int(lens.sum())).astype(np.int64)
    y = pm(torch.tensor(idx), torch.tensor(off))
    gy = rs.randn(*y.shape).astype(np.float32)
    y.backward(torch.tensor(gy))
    ge = pm.embs.weight.grad.coalesce()
    out.update({"pr_w": pm.embs.weight.detach().numpy(), "pr_proj": pm.proj.weight.detach().numpy(), "pr_idx": idx,
                "pr_off": off, "pr_y": y.detach().numpy(), "pr_gy": gy, "pr_ge_idx": ge.indices().numpy()[0],
                "pr_ge_val": ge.values().numpy(), "pr_gproj": pm.proj.weight.grad.numpy()})
    # dimension solver
    n = torch.tensor([39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346, 10])
    for a_, alpha in enumerate((0.0, 0.25, 0.5)):
        out[f"md_d0_{a_}"] = md.md_solver(n.clone(), alpha, d0=128, round_dim=True).numpy()
        out[f"md_raw_{a_}"] = md.md_solver(n.clone(), alpha, d0=128, round_dim=False).numpy()
    out["md_n"] = n.numpy()
    out["md_budget"] = md.md_solver(n.clone(), 0.3, B=1.0e8, round_dim=False).numpy()
    np.savez_compressed(os.path.join(OUT, "tricks.npz"), **out)


def load_T_model():
    """T/dlrm_s_pytorch (the training/bfloat16 variant of the model) + its data module, imported as modules: the
    absent third-party imports at T/dlrm_s_pytorch.py:90,102-103 and T/mlperf_logger.py:26-27 are stubbed."""
    sys.dont_write_bytecode = True
    _stub("onnx")
    core = _stub("intel_pytorch_extension.core")
    _stub("intel_pytorch_extension", core=core, is_available=lambda: False)
    consts = _stub("mlperf_logging.mllog.constants")
    mllog = _stub("mlperf_logging.mllog", constants=consts, get_mllogger=lambda: types.SimpleNamespace())
    _stub("mlperf_logging", mllog=mllog)
    for n in _SHARED_NAMES + ["dlrm_s_pytorch"]:
        sys.modules.pop(n, None)
    sys.path.insert(0, T_DIR)
    try:
        TM = importlib.import_module("dlrm_s_pytorch")
        dp = importlib.import_module("dlrm_data_pytorch")
    finally:
        sys.path.remove(T_DIR)
    for n in _SHARED_NAMES + ["dlrm_s_pytorch"]:
        sys.modules.pop(n, None)
    return TM, dp


def gen_T_smoke():
    """The reference's OWN test configurations (T/test/dlrm_s_test.sh:24 and :26 -- the only tests it has for this
    path; they diff PyTorch against Caffe2 at run time and store no values).  Same sequence as the script's
    __main__ for ``--mini-batch-size=2 --data-size=4|5 --nepochs=1|3 --arch-interaction-op=dot --learning-rate=0.1``
    with every other flag at its default (tables 4-3-2, D = 2, bottom 4-3-2, top 4-2-1 -> 8-4-2-1, MSE loss,
    seed 123, up to 10 indices per lookup): seeds (T/dlrm_s_pytorch.py:743-746), random data loader (:811), model
    (:945), MSELoss (:1001), SGD (:1018), loop (:1184-1256).  Expected loss traces (SURVEY.md section 8c /
    BASELINE.md): 0.028790, 0.156787 and the nine values of the 3-epoch run -- asserted here."""
    TM, dp = load_T_model()
    out = {}
    want = {"c2": [0.028790, 0.156787],
            "c4": [0.028790, 0.156787, 0.183348, 0.022413, 0.134285, 0.165428, 0.016494, 0.111174, 0.137763]}
    for tag, data_size, nepochs in (("c2", 4, 1), ("c4", 5, 3)):
        args = types.SimpleNamespace(data_size=data_size, num_batches=0, mini_batch_size=2, num_indices_per_lookup=10,
                                     num_indices_per_lookup_fixed=False, round_targets=False, data_generation="random",
                                     data_trace_file="./input/dist_emb_j.log", data_trace_enable_padding=False,
                                     numpy_rand_seed=123, num_workers=0)
        np.random.seed(123)
        torch.manual_seed(123)
        ln_emb, ln_bot, m_spa = np.array([4, 3, 2]), np.array([4, 3, 2]), 2
        _, train_ld = dp.make_random_data_and_loader(args, ln_emb, 4)
        ln_top = np.array([4 * 3 // 2 + 2, 4, 2, 1])
        net = TM.DLRM_Net(m_spa, ln_emb, ln_bot, ln_top, arch_interaction_op="dot", arch_interaction_itself=False,
                          sigmoid_bot=-1, sigmoid_top=ln_top.size - 2, sync_dense_params=True, loss_threshold=0.0,
                          ndevices=-1)
        params0 = [p.detach().numpy().copy() for p in net.parameters()]
        names = [n for n, _ in net.named_parameters()]
        loss_fn = torch.nn.MSELoss(reduction="mean")
        opt = torch.optim.SGD(net.parameters(), lr=0.1)
        losses, step = [], 0
        for _ep in range(nepochs):
            for (X, lS_o, lS_i, T) in train_ld:
                E = loss_fn(net(X, lS_o, lS_i), T)
                opt.zero_grad()
                E.backward()
                opt.step()
                losses.append(float(E.detach()))
                out[f"{tag}_X{step}"], out[f"{tag}_T{step}"] = X.numpy(), T.numpy()
                for t in range(3):
                    out[f"{tag}_o{step}_{t}"], out[f"{tag}_i{step}_{t}"] = lS_o[t].numpy(), lS_i[t].numpy()
                step += 1
        assert [round(v, 6) for v in losses] == want[tag], (tag, losses)
        out[f"{tag}_losses"], out[f"{tag}_steps"] = np.array(losses, dtype=np.float64), np.int64(step)
        out[f"{tag}_param_names"] = np.array(names)
        for k, (a, b) in enumerate(zip(params0, net.parameters())):
            out[f"{tag}_init{k}"], out[f"{tag}_final{k}"] = a, b.detach().numpy().copy()
    out["ln_emb"], out["ln_bot"], out["ln_top"] = np.array([4, 3, 2]), np.array([4, 3, 2]), np.array([8, 4, 2, 1])
    np.savez_compressed(os.path.join(OUT, "t_smoke.npz"), **out)


def gen_T_options():
    """Model options of the training variant that the product variant lacks, from T/dlrm_s_pytorch.py itself:
    --arch-interaction-itself (T/...:428, pairs j <= i) forward + autograd backward, --arch-interaction-op=cat
    (:434-436) and the wbce loss (:1004-1006,1066-1075)."""
    TM, _ = load_T_model()
    net = TM.DLRM_Net.__new__(TM.DLRM_Net)
    torch.nn.Module.__init__(net)
    net.bf16 = False
    out = {}
    rs = np.random.RandomState(21)
    for tag, (B, F, D) in {"a": (5, 4, 8), "b": (3, 27, 16), "c": (2, 27, 128)}.items():
        net.arch_interaction_op, net.arch_interaction_itself = "dot", True
        x = torch.tensor(rs.randn(B, D).astype(np.float32), requires_grad=True)
        ly = [torch.tensor(rs.randn(B, D).astype(np.float32), requires_grad=True) for _ in range(F - 1)]
        R = net.interact_features(x, ly)
        dR = torch.tensor(rs.randn(*R.shape).astype(np.float32))
        R.backward(dR)
        out[f"{tag}_x"], out[f"{tag}_ly"] = x.detach().numpy(), np.stack([t.detach().numpy() for t in ly])
        out[f"{tag}_R"], out[f"{tag}_dR"] = R.detach().numpy(), dR.numpy()
        out[f"{tag}_dx"], out[f"{tag}_dly"] = x.grad.numpy(), np.stack([t.grad.numpy() for t in ly])
        net.arch_interaction_op = "cat"
        out[f"{tag}_cat"] = net.interact_features(x.detach(), [t.detach() for t in ly]).numpy()
    # wbce: per-sample BCE weighted by loss_ws[target], then the mean (T/dlrm_s_pytorch.py:1066-1075)
    Z = torch.tensor(rs.rand(16, 1).astype(np.float32) * 0.98 + 0.01)
    Tt = torch.tensor(np.round(rs.rand(16, 1)).astype(np.float32))
    loss_ws = torch.tensor(np.fromstring("1.0-3.0", dtype=float, sep="-"))
    loss_fn = torch.nn.BCELoss(reduction="none")
    ws_ = loss_ws[Tt.data.view(-1).long()].view_as(Tt)
    out["wbce_Z"], out["wbce_T"], out["wbce_ws"] = Z.numpy(), Tt.numpy(), loss_ws.numpy()
    out["wbce_loss"] = np.float64((ws_ * loss_fn(Z, Tt)).mean().item())
    np.savez_compressed(os.path.join(OUT, "t_options.npz"), **out)


def gen_trace_synth():
    """The trace-driven index generator (``--data-generation=synthetic``), run through the reference's own functions:
    ``trace_profile`` (T/dlrm_data_pytorch.py:954) on a seeded skewed id trace per table, the distribution built as
    the module's ``__main__`` does (:1088-1112; restated here, a script body cannot be imported), the three-line
    files (:1045), then ``trace_generate_lru`` (:902) and ``generate_synthetic_input_batch`` (:807) under
    ``np.random.seed``.  The file contents are stored in the fixture (the test writes them back out)."""
    import collections
    import tempfile
    G = load_T_generator()
    rs = np.random.RandomState(99)
    ln_emb = [50, 400, 7]
    out = {"ln_emb": np.array(ln_emb)}
    tmp = tempfile.mkdtemp()
    for t, rows in enumerate(ln_emb):
        trace = np.minimum((rs.pareto(1.2, size=600) * 3).astype(np.int64), rows - 1)
        out[f"trace{t}"] = trace
        for pad in (False, True):
            _, sds, lines = G.trace_profile([np.uint64(x) for x in trace], pad)
            sds.reverse()
            lines.reverse()
            n = len(sds)
            dc = sorted(collections.Counter(sds).items(), key=lambda kv: kv[0])
            list_sd = [int(k) for k, _ in dc]
            cumm = []
            for i, (_, k) in enumerate(dc):
                cumm.append(k / float(n) if i == 0 else cumm[i - 1] + k / float(n))
            tag = f"{t}_{int(pad)}"
            out[f"lines{tag}"] = np.array([int(x) for x in lines], dtype=np.int64)
            out[f"dist{tag}"] = np.array(list_sd, dtype=np.int64)
            out[f"cumm{tag}"] = np.array(cumm, dtype=np.float64)
            G.write_dist_to_file(os.path.join(tmp, f"dist{int(pad)}_emb_{t}.log"), [int(x) for x in lines], list_sd, cumm)
            np.random.seed(7 + t)
            la, sd, cs = G.read_dist_from_file(os.path.join(tmp, f"dist{int(pad)}_emb_{t}.log"))
            out[f"lru{tag}"] = np.array(G.trace_generate_lru(la, sd, cs, 300, pad), dtype=np.uint64)
    for tag, fixed, pad in (("ragged", False, False), ("fixed", True, False), ("ragged_pad", False, True)):
        np.random.seed(123)
        X, lS_o, lS_i = G.generate_synthetic_input_batch(4, ln_emb, 6, 5, fixed, os.path.join(tmp, f"dist{int(pad)}_emb_j.log"), pad)
        out[f"{tag}_X"] = X.numpy()
        for t in range(len(ln_emb)):
            out[f"{tag}_o{t}"] = lS_o[t].numpy()
            out[f"{tag}_i{t}"] = lS_i[t].numpy()
    np.savez_compressed(os.path.join(OUT, "trace_synth.npz"), **out)


def gen_collate_criteo():
    """The per-sample Criteo collate (T/dlrm_data_pytorch.py:339-352) run on seeded integer samples."""
    G = load_T_generator()
    rs = np.random.RandomState(17)
    samples = [(rs.randint(0, 5000, size=13).astype(np.int32), rs.randint(0, 10 ** 6, size=26).astype(np.int32),
                np.int32(rs.randint(0, 2))) for _ in range(9)]
    X, lS_o, lS_i, T = G.collate_wrapper_criteo([(a.tolist(), b.tolist(), int(c)) for a, b, c in samples])
    np.savez_compressed(os.path.join(OUT, "collate_criteo.npz"), x_int=np.stack([s_[0] for s_ in samples]),
                        x_cat=np.stack([s_[1] for s_ in samples]), y=np.array([s_[2] for s_ in samples]),
                        X=X.numpy(), lS_o=lS_o.numpy(), lS_i=lS_i.numpy(), T=T.numpy())


def main():
    os.makedirs(OUT, exist_ok=True)
    G = load_T_generator()
    M, X = load_P()
    gen_interaction(M)
    gen_index_kats(X)
    gen_lr_policy(M)
    gen_small_model(M, G)
    gen_kaggle_c1(M, G)
    gen_criteo_transform(load_T_terabyte())
    gen_int8_tables()
    gen_tricks()
    gen_T_smoke()
    gen_T_options()
    gen_trace_synth()
    gen_collate_criteo()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
