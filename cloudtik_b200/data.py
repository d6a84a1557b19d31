"""Host-side producers of the batch format the DLRM step consumes.

Mirrors of the reference's random-data path, ``T/dlrm_data_pytorch.py``: ``RandomDataset`` (``:541-640``),
``generate_uniform_input_batch`` (``:759-803``), ``generate_random_output_batch`` (``:748-755``),
``collate_wrapper_random`` (``:646-652``) and ``make_random_data_and_loader`` (``:655-682``).  Same argument
meaning, same return format ``(X, lS_o, lS_i, T)`` and -- because the reference draws everything from the legacy
numpy stream in a fixed order -- the SAME VALUES for the same seed: a run started with the reference's flags sees
the reference's batches (pinned by ``tests/golden/t_smoke.npz`` and ``small_model.npz``).

This is plain numpy on the host (index arithmetic, no table data): the tensors it returns are what the caller
copies to the device.  The stream lives in a private ``RandomState`` instead of the process-global one the
reference reseeds; ``RandomState(seed)`` and ``np.random.seed(seed)`` produce the same sequence.
"""
from __future__ import annotations

import numpy as np
import torch


def generate_random_output_batch(n, num_targets, round_targets=False, ra=np.random):
    """Targets: U[0,1) as float32, rounded to 0/1 on request."""
    p = ra.rand(n, num_targets).astype(np.float32)
    return torch.tensor(np.round(p).astype(np.float32) if round_targets else p)


def generate_uniform_input_batch(m_den, ln_emb, n, num_indices_per_lookup, num_indices_per_lookup_fixed, ra=np.random):
    """Dense ``[n, m_den]`` U[0,1) and, per table, n bags: a bag draws its size (fixed, or
    ``round(max(1, u * min(rows, P)))``), then that many ``round(u * (rows - 1))`` ids, sorted and de-duplicated
    (so bags are ragged).  Draw order = the reference's: dense block first, then table by table, sample by sample."""
    X = torch.tensor(ra.rand(n, m_den).astype(np.float32))
    lS_o, lS_i = [], []
    for rows in ln_emb:
        rows = int(rows)
        starts, ids = np.empty(n, dtype=np.int64), []
        filled = 0
        for b in range(n):
            if num_indices_per_lookup_fixed:
                k = int(num_indices_per_lookup)
            else:
                u = float(ra.random_sample(1)[0])
                k = int(np.round(max(1.0, u * min(rows, num_indices_per_lookup))))
            bag = np.unique(np.round(ra.random_sample(k) * (rows - 1)).astype(np.int64))
            starts[b] = filled
            ids.append(bag)
            filled += bag.size
        lS_o.append(torch.tensor(starts))
        lS_i.append(torch.tensor(np.concatenate(ids) if ids else np.empty(0, dtype=np.int64)))
    return X, lS_o, lS_i


class RandomDataset(torch.utils.data.Dataset):
    """One item = one whole mini-batch ``(X, lS_o, lS_i, T)``; with ``reset_seed_on_access`` the stream restarts at
    item 0, so every epoch sees the same batches (the reference's behaviour for ``--data-generation=random``)."""

    def __init__(self, m_den, ln_emb, data_size, num_batches, mini_batch_size, num_indices_per_lookup,
                 num_indices_per_lookup_fixed, num_targets=1, round_targets=False, data_generation="random",
                 trace_file="", enable_padding=False, reset_seed_on_access=False, rand_seed=0, world_size=1):
        if data_generation != "random":
            raise NotImplementedError("only --data-generation=random (no trace files in this build)")
        nbatches = int(np.ceil((data_size * 1.0) / mini_batch_size))
        if num_batches != 0:
            nbatches = num_batches
            data_size = nbatches * mini_batch_size
        self.m_den, self.ln_emb = m_den, ln_emb
        self.data_size, self.num_batches, self.mini_batch_size = data_size, nbatches, mini_batch_size
        self.num_indices_per_lookup = num_indices_per_lookup
        self.num_indices_per_lookup_fixed = num_indices_per_lookup_fixed
        self.num_targets, self.round_targets = num_targets, round_targets
        self.reset_seed_on_access, self.rand_seed = reset_seed_on_access, rand_seed
        self.world_size = world_size
        self.ra = np.random.RandomState(rand_seed)

    def __len__(self):
        return self.num_batches

    def __getitem__(self, index):
        if isinstance(index, slice):
            return [self[i] for i in range(index.start or 0, index.stop or len(self), index.step or 1)]
        if self.reset_seed_on_access and index == 0:
            self.ra = np.random.RandomState(self.rand_seed)
        n = min(self.mini_batch_size, self.data_size - index * self.mini_batch_size)
        if self.world_size > 1:
            n = n // self.world_size
        X, lS_o, lS_i = generate_uniform_input_batch(self.m_den, self.ln_emb, n, self.num_indices_per_lookup,
                                                     self.num_indices_per_lookup_fixed, ra=self.ra)
        T = generate_random_output_batch(n, self.num_targets, self.round_targets, ra=self.ra)
        return X, lS_o, lS_i, T


def collate_wrapper_random(list_of_tuples):
    """DataLoader(batch_size=1) collate: unwrap the single pre-built batch; offsets become one ``[T, B]`` tensor."""
    X, lS_o, lS_i, T = list_of_tuples[0]
    return X, torch.stack(lS_o), lS_i, T


def make_random_data_and_loader(args, ln_emb, m_den, world_size=1):
    data = RandomDataset(m_den, ln_emb, args.data_size, args.num_batches, args.mini_batch_size,
                         args.num_indices_per_lookup, args.num_indices_per_lookup_fixed, 1, args.round_targets,
                         args.data_generation, reset_seed_on_access=True, rand_seed=args.numpy_rand_seed,
                         world_size=world_size)
    loader = torch.utils.data.DataLoader(data, batch_size=1, shuffle=False, num_workers=0,
                                         collate_fn=collate_wrapper_random, pin_memory=False, drop_last=False)
    return data, loader
