"""Inference / evaluation path on the device (SURVEY.md section 8f-3 and the eval half of 8f-4).

Mirror of the reference's ``inference(args, dlrm, best_acc_test, best_auc_test, test_ld)``
(``P/dlrm_s_pytorch.py:478-639``): same arguments, same progress / metric lines, same returned
``(model_metrics_dict, is_best)``.  What differs is where the work happens:

* the forward is the forward-only half of the training kernels (``b200_emb_fwd`` + ``b200_interact_fwd``
  under ``torch.no_grad()``), no tracing / freezing step;
* scores and targets never visit the host batch by batch (the reference does ``Z_test.detach().cpu()`` per
  batch, ``:543-553``): accuracy is counted on the device, and the MLPerf metrics (``:563-584``: recall,
  precision, f1, average precision, ROC AUC, accuracy via sklearn) are computed from ONE device sort of the
  scores -- ``b200_sort_dedup`` on an order-preserving integer image of the fp32 scores; its run table is
  exactly the list of distinct thresholds with their tie groups.
"""
from __future__ import annotations

import time

import numpy as np
import torch

from . import ops


def _score_keys(scores: torch.Tensor) -> torch.Tensor:
    """fp32 scores -> uint32 keys (as int32 bit patterns) whose unsigned order is the float order."""
    bits = scores.detach().float().contiguous().view(-1).view(torch.int32)
    return torch.where(bits < 0, ~bits, bits ^ (-2 ** 31))


def threshold_table(scores: torch.Tensor, targets: torch.Tensor):
    """Distinct score values in ASCENDING order with the number of samples / positives at each:
    (n_in, pos_in) int64 tensors, one entry per distinct score.  NaN scores are not allowed."""
    n = scores.numel()
    if n == 0:
        raise ValueError("no samples")
    keys = _score_keys(scores)
    vals = torch.arange(n, dtype=torch.int32, device=scores.device)
    _, sv, _, seg, _ = ops.sort_dedup(keys, vals, 1 << 32)
    t_sorted = (targets.detach().reshape(-1)[sv.long()] > 0.5).to(torch.int64)
    cpos = torch.cumsum(t_sorted, 0)
    seg = seg.long()
    pos_end = cpos[seg[1:] - 1]
    pos_in = pos_end - torch.cat([pos_end.new_zeros(1), pos_end[:-1]])
    n_in = seg[1:] - seg[:-1]
    return n_in, pos_in


def roc_auc_score(targets: torch.Tensor, scores: torch.Tensor) -> float:
    """sklearn.metrics.roc_auc_score for binary targets, on the device (ties share a threshold: trapezoids)."""
    n_in, pos_in = threshold_table(scores, targets)
    neg_in = n_in - pos_in
    P, N = int(pos_in.sum()), int(neg_in.sum())
    if P == 0 or N == 0:
        raise ValueError("Only one class present in y_true. ROC AUC score is not defined in that case.")
    neg_before = torch.cumsum(neg_in, 0) - neg_in
    num = (pos_in.double() * (neg_before.double() + 0.5 * neg_in.double())).sum()
    return float(num / (float(P) * float(N)))


def average_precision_score(targets: torch.Tensor, scores: torch.Tensor) -> float:
    """sklearn.metrics.average_precision_score (binary): sum over distinct thresholds, highest first, of
    (recall_k - recall_{k-1}) * precision_k."""
    n_in, pos_in = threshold_table(scores, targets)
    n_in, pos_in = n_in.flip(0), pos_in.flip(0)
    tp = torch.cumsum(pos_in, 0).double()
    seen = torch.cumsum(n_in, 0).double()
    P = float(tp[-1])
    if P == 0:
        return 0.0
    return float(((pos_in.double() / P) * (tp / seen)).sum())


def binary_metrics(targets: torch.Tensor, scores: torch.Tensor) -> dict:
    """The six metrics of ``P/dlrm_s_pytorch.py:563-584`` from device tensors."""
    t = targets.detach().reshape(-1) > 0.5
    p = torch.round(scores.detach().float().reshape(-1)) > 0.5
    tp = int((t & p).sum())
    fp = int((~t & p).sum())
    fn = int((t & ~p).sum())
    n = t.numel()
    recall = tp / (tp + fn) if tp + fn else 0.0
    precision = tp / (tp + fp) if tp + fp else 0.0
    f1 = 2 * precision * recall / (precision + recall) if precision + recall else 0.0
    return {"recall": recall, "precision": precision, "f1": f1,
            "ap": average_precision_score(targets, scores), "roc_auc": roc_auc_score(targets, scores),
            "accuracy": (n - fp - fn) / n}


def _unpack(batch, device):
    X, lS_o, lS_i, T = batch[0], batch[1], batch[2], batch[3]
    to = (lambda v: [u.to(device, non_blocking=True) for u in v] if isinstance(v, (list, tuple))
          else v.to(device, non_blocking=True))
    return to(X), to(lS_o), to(lS_i), to(T)


def inference(args, dlrm, best_acc_test, best_auc_test, test_ld, device=None, nbatches=0, nbatches_test=0):
    """``P/dlrm_s_pytorch.py:478-639``.  ``args`` needs ``print_auc, inference_only, print_freq, nepochs,
    mlperf_auc_threshold`` (missing ones default to the reference's defaults)."""
    g = lambda k, d: getattr(args, k, d)  # noqa: E731
    print_auc, inference_only = g("print_auc", False), g("inference_only", False)
    print_freq = g("print_freq", 1)
    device = device or next(dlrm.parameters()).device
    was_training = dlrm.training
    dlrm.eval()
    scores, targets = [], []
    correct = torch.zeros((), dtype=torch.int64, device=device)
    test_samp = 0
    total_time, total_iter = 0.0, 0
    with torch.no_grad():
        for i, batch in enumerate(test_ld):
            if inference_only and nbatches > 0 and i >= nbatches:
                break
            X, lS_o, lS_i, T = _unpack(batch, device)
            torch.cuda.synchronize(device)
            start = time.time()
            if isinstance(lS_i, (list, tuple)) or lS_i.dim() == 2:
                Z = dlrm(X, list(lS_o) if not isinstance(lS_o, list) else lS_o,
                         list(lS_i) if not isinstance(lS_i, list) else lS_i)
            else:  # merged form: (X, offsets, indices, T) as CriteoBinDataset(merged=True) yields
                Z = dlrm(X, lS_i, lS_o, None)
            torch.cuda.synchronize(device)
            total_time += time.time() - start
            total_iter += 1
            if inference_only and ((i + 1) % print_freq == 0 or i + 1 == len(test_ld)):
                print("Finished {} it {}/{}, {:.2f} ms/it,".format("inference", i + 1, len(test_ld),
                                                                   1000.0 * total_time / total_iter), flush=True)
                total_time, total_iter = 0.0, 0
            if print_auc:
                scores.append(Z.detach().float().reshape(-1))
                targets.append(T.detach().float().reshape(-1))
            elif not inference_only:
                correct += (torch.round(Z.detach().float()) == T.float()).sum()
                test_samp += T.shape[0]
    dlrm.train(was_training)

    validation_results = None
    acc_test = None
    if print_auc:
        validation_results = binary_metrics(torch.cat(targets), torch.cat(scores))
        acc_test = validation_results["accuracy"]
    elif not inference_only:
        acc_test = int(correct.item()) / max(test_samp, 1)

    model_metrics_dict = {"nepochs": g("nepochs", 1), "nbatches": nbatches, "nbatches_test": nbatches_test}
    if not inference_only:
        model_metrics_dict["test_acc"] = acc_test
    is_best = False
    if print_auc:
        is_best = validation_results["roc_auc"] > best_auc_test
        if is_best:
            best_auc_test = validation_results["roc_auc"]
            model_metrics_dict["test_auc"] = best_auc_test
        print("recall {:.4f}, precision {:.4f},".format(validation_results["recall"], validation_results["precision"])
              + " f1 {:.4f}, ap {:.4f},".format(validation_results["f1"], validation_results["ap"])
              + " auc {:.4f}, best auc {:.4f},".format(validation_results["roc_auc"], best_auc_test)
              + " accuracy {:3.3f} %, best accuracy {:3.3f} %".format(validation_results["accuracy"] * 100,
                                                                    best_acc_test * 100), flush=True)
        print("Accuracy: {:.34} ".format(validation_results["roc_auc"]))
    elif not inference_only:
        is_best = acc_test > best_acc_test
        if is_best:
            best_acc_test = acc_test
        print(" accuracy {:3.3f} %, best {:3.3f} %".format(acc_test * 100, best_acc_test * 100), flush=True)
    if not inference_only:
        return model_metrics_dict, is_best
    return None


class GraphedInference:
    """Forward-only step replayed as one CUDA graph: the serving-throughput counterpart of the reference's
    ``run_throughput_benchmark`` (``P/dlrm_s_pytorch.py:456-476``).  ``example`` = (X, indices, offsets)."""

    def __init__(self, net, example, warmup: int = 3):
        self.net = net
        self.static = [torch.empty_like(t) for t in example]
        for d, s in zip(self.static, example):
            d.copy_(s)
        net.eval()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(warmup):
                self._fwd()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.out = self._fwd()
        torch.cuda.synchronize()

    def _fwd(self):
        X, idx, off = self.static
        return self.net(X, idx, off, None).float()

    def __call__(self, X, indices, offsets):
        for d, s in zip(self.static, (X, indices, offsets)):
            if d.data_ptr() != s.data_ptr():
                d.copy_(s, non_blocking=True)
        self.graph.replay()
        return self.out


def sklearn_reference_metrics(targets: np.ndarray, scores: np.ndarray) -> dict:
    """The reference's own metric table (``P/dlrm_s_pytorch.py:563-584``), for tests."""
    import sklearn.metrics as m
    yp = np.round(scores)
    return {"recall": m.recall_score(targets, yp), "precision": m.precision_score(targets, yp),
            "f1": m.f1_score(targets, yp), "ap": m.average_precision_score(targets, scores),
            "roc_auc": m.roc_auc_score(targets, scores), "accuracy": m.accuracy_score(targets, yp)}
