"""Criteo-Terabyte binary input on the device.

Host-side mirror of the reference's binary loader interface, ``T/data_loader_terabyte.py``:

* ``_transform_features(x_int_batch, x_cat_batch, y_batch, max_ind_range, flag_input_torch_tensor)``
  (``:83-102``) keeps its name, argument meaning and return tuple ``(X, lS_o, lS_i, Y)``; the inputs are
  views of ONE device tensor of raw int32 records, the outputs come from ONE kernel
  (``b200_criteo_transform``);
* ``CriteoBinDataset`` (``:207-271``): same constructor arguments, ``__len__`` / ``__getitem__``; it reads
  the same file format (int32 records ``label, 13 counts, 26 ids`` written by ``numpy_to_binary`` ``:274-306``)
  into pinned host memory, copies the raw records (160 B per sample instead of the 472 B of transformed
  tensors) and transforms them on the GPU.  Ranks of a multi-process job read the GLOBAL batch (every rank
  needs all samples' ids for its tables) and keep the dense part of their own slice, which is what the
  table-wise model-parallel forward consumes.

No CPU path: without the CUDA library the import of ``cloudtik_b200.ops`` fails.
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch

from . import ext_dist, ops


def transform_records(records: torch.Tensor, max_ind_range: int = -1, n_dense: int = 13, **kw):
    """Merged-embedding form: ``(X [B, 13], Y [B, 1], indices [T*B], offsets [T*B+1])`` from raw records."""
    return ops.criteo_transform(records, max(int(max_ind_range), 0), n_dense=n_dense, **kw)


def _transform_features(x_int_batch, x_cat_batch, y_batch, max_ind_range, flag_input_torch_tensor=False):
    """Reference signature (``T/data_loader_terabyte.py:83-102``).  The three inputs must be the column
    views ``rec[:, 1:14]``, ``rec[:, 14:]``, ``rec[:, 0]`` of one contiguous int32 CUDA tensor ``rec`` -- the
    way ``CriteoBinDataset.__getitem__`` passes them (``:265-271``)."""
    base = x_cat_batch._base if x_cat_batch._base is not None else x_cat_batch
    n_dense = x_int_batch.shape[1]
    ok = (base.dim() == 2 and base.dtype == torch.int32 and base.is_cuda and base.is_contiguous()
          and base.shape[1] == 1 + n_dense + x_cat_batch.shape[1]
          and x_int_batch.data_ptr() == base.data_ptr() + 4 and y_batch.data_ptr() == base.data_ptr()
          and x_cat_batch.data_ptr() == base.data_ptr() + 4 * (1 + n_dense))
    if not ok:
        raise RuntimeError("_transform_features: pass column views of one contiguous int32 CUDA record tensor")
    X, Y, ind, off = transform_records(base, max_ind_range, n_dense=n_dense)
    B, T = base.shape[0], x_cat_batch.shape[1]
    lS_i = ind.view(T, B)
    lS_o = (off[:-1].view(T, B) - (torch.arange(T, device=off.device) * B).view(-1, 1))
    return X, lS_o, lS_i, Y


class CriteoBinDataset(torch.utils.data.Dataset):
    """Binary version of the Criteo dataset, decoded on the GPU (``T/data_loader_terabyte.py:207-271``)."""

    def __init__(self, data_file, counts_file=None, batch_size=1, max_ind_range=-1, bytes_per_feature=4,
                 device=None, merged=False):
        self.tar_fea = 1
        self.den_fea = 13
        self.spa_fea = 26
        self.tad_fea = self.tar_fea + self.den_fea
        self.tot_fea = self.tad_fea + self.spa_fea
        if bytes_per_feature != 4:
            raise NotImplementedError("records are int32 (numpy_to_binary writes int32)")
        self.batch_size = batch_size
        self.max_ind_range = max_ind_range
        self.bytes_per_sample = bytes_per_feature * self.tot_fea
        self.bytes_per_batch = self.bytes_per_sample * batch_size
        data_file_size = os.path.getsize(data_file)
        self.num_batches = math.ceil(data_file_size / self.bytes_per_batch)
        self.num_samples = data_file_size // self.bytes_per_sample
        self.file = open(data_file, "rb")
        self.counts = None
        if counts_file is not None:
            with np.load(counts_file) as data:
                self.counts = data["counts"]
        self.m_den = 13
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.merged = merged
        self._pinned = torch.empty(batch_size * self.tot_fea, dtype=torch.int32).pin_memory()
        self._copied = None  # event: the last H2D copy out of the pinned buffer

    def __len__(self):
        return self.num_batches

    def read_records(self, idx) -> torch.Tensor:
        """Raw records of batch ``idx`` on the device, int32 [n, 40] (last batch may be short)."""
        self.file.seek(idx * self.bytes_per_batch, 0)
        raw = self.file.read(self.bytes_per_batch)
        n = len(raw) // self.bytes_per_sample
        host = self._pinned[:n * self.tot_fea]
        if self._copied is not None:
            self._copied.synchronize()  # the previous batch has left the pinned buffer
        host.numpy()[:] = np.frombuffer(raw, dtype=np.int32, count=n * self.tot_fea)
        rec = host.view(n, self.tot_fea).to(self.device, non_blocking=True)
        self._copied = torch.cuda.Event()
        self._copied.record()
        return rec

    def __getitem__(self, idx):
        rec = self.read_records(idx)
        world, rank = max(ext_dist.my_size, 1), max(ext_dist.my_rank, 0)
        n = rec.shape[0]
        if self.merged or world > 1:
            # global-batch ids of this rank's tables, dense part / labels of this rank's samples
            sl = ext_dist.get_my_slice(self.spa_fea) if world > 1 else slice(0, self.spa_fea, 1)
            ln = n // world
            X, Y, ind, off = transform_records(rec, self.max_ind_range, n_dense=self.den_fea, tab_begin=sl.start,
                                               tab_count=sl.stop - sl.start, row_begin=rank * ln, row_count=ln)
            return X, off, ind, Y
        return _transform_features(x_int_batch=rec[:, 1:14], x_cat_batch=rec[:, 14:], y_batch=rec[:, 0],
                                   max_ind_range=self.max_ind_range, flag_input_torch_tensor=True)


def numpy_to_binary(y, x_int, x_cat, output_file_path):
    """Write records in the format ``CriteoBinDataset`` reads (``T/data_loader_terabyte.py:274-306``):
    int32 ``[label, 13 counts, 26 ids]`` per sample."""
    rec = np.concatenate([np.asarray(y).reshape(-1, 1), np.asarray(x_int), np.asarray(x_cat)], axis=1).astype(np.int32)
    with open(output_file_path, "wb") as f:
        f.write(rec.tobytes())
    return rec
