"""Host-side producers of the batch format the DLRM step consumes.

Mirrors of the reference's random-data path, ``T/dlrm_data_pytorch.py``: ``RandomDataset`` (``:541-640``),
``generate_uniform_input_batch`` (``:759-803``), ``generate_random_output_batch`` (``:748-755``),
``collate_wrapper_random`` (``:646-652``), ``collate_wrapper_criteo`` (``:339-352``) and
``make_random_data_and_loader`` (``:655-682``).  Same argument
meaning, same return format ``(X, lS_o, lS_i, T)`` and -- because the reference draws everything from the legacy
numpy stream in a fixed order -- the SAME VALUES for the same seed: a run started with the reference's flags sees
the reference's batches (pinned by ``tests/golden/t_smoke.npz`` and ``small_model.npz``).

``--data-generation=synthetic`` (``generate_synthetic_input_batch`` ``:807-878``) is here too: ids drawn by replaying
an access trace's stack-distance distribution through an LRU stack (``trace_profile`` ``:954-1000``, the
distribution built as the module's ``__main__`` does ``:1088-1112``, ``trace_generate_lru`` ``:902-927``,
``generate_stack_distance`` ``:881-895``, the three-line distribution file ``:1030-1057``).  Same draws in the same
order, so the same bags for the same seed and distribution files (pinned by ``tests/golden/trace_synth.npz``, written
by the reference's functions).

This is plain numpy on the host (index arithmetic, no table data): the tensors it returns are what the caller
copies to the device.  The stream lives in a private ``RandomState`` instead of the process-global one the
reference reseeds; ``RandomState(seed)`` and ``np.random.seed(seed)`` produce the same sequence.
"""
from __future__ import annotations

import bisect
from collections import Counter

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


class StackDistanceModel:
    """An access pattern as (lines in first-touch order, stack distances seen, their cumulative probabilities):
    what the reference keeps in a ``dist_emb_<table>.log`` file.  Distance 0 = a line not touched before; distance
    d > 0 = the line d places below the top of the LRU stack (the reference counts the top as 1)."""

    def __init__(self, lines, distances, cumulative):
        self.lines = [int(x) for x in lines]
        self.distances = [int(x) for x in distances]
        self.cumulative = np.asarray([float(x) for x in cumulative], dtype=np.float64)

    @classmethod
    def from_trace(cls, trace, enable_padding=False):
        """Profile an id trace.  One pass with the stack kept oldest-first: a hit at position i of l entries has
        distance l - i and moves to the top; a miss has distance 0 and is appended to the line list."""
        stack, lines, dists = [], [], []
        for x in trace:
            r = int(x)
            try:
                i = stack.index(r)
                dists.append(len(stack) - i)
                stack.append(stack.pop(i))
            except ValueError:
                dists.append(0)
                lines.append(r)
                stack.append(r)
        if enable_padding:  # extra zero distances so that new lines keep appearing (reference: l / max distance)
            dists += [0] * int(np.ceil(len(dists) / max(dists)))
        total = float(len(dists))
        hist = sorted(Counter(dists).items())
        cumulative, acc = [], 0.0
        for k, (_, c) in enumerate(hist):
            acc = c / total if k == 0 else acc + c / total
            cumulative.append(acc)
        return cls(lines, [d for d, _ in hist], cumulative)

    @classmethod
    def read(cls, path):
        """Three comma-separated lines: first-touch lines, distances, cumulative probabilities."""
        with open(path, "r") as f:
            rows = f.read().splitlines()
        return cls(*[[tok for tok in row.split(", ")] for row in rows[:3]])

    def write(self, path):
        with open(path, "w") as f:
            for row in (self.lines, self.distances, [float(c) for c in self.cumulative]):
                f.write(str(list(row))[1:-1] + "\n")

    def _draw_distance(self, seen, enable_padding, ra):
        u = ra.rand(1)
        if seen < self.distances[-1]:
            # only distances reachable with the lines seen so far: scale u into their share of the distribution
            u = u * self.cumulative[bisect.bisect(self.distances, seen) - 1]
        elif enable_padding:
            lo = self.cumulative[0]
            u = (1.0 - lo) * u + lo
        j = int(np.searchsorted(self.cumulative, u[0], side="left"))
        return self.distances[j]

    def generate_lru(self, length, enable_padding=False, ra=np.random):
        """``length`` ids: a fresh copy of the line list is the LRU stack (front = next unseen line, back = most
        recent); a distance-0 draw takes the front, a distance-d draw the entry d from the back; either moves to the
        back.  One uniform draw per id."""
        stack = list(self.lines)
        depth, seen, out = len(stack), 0, np.empty(int(length), dtype=np.uint64)
        for k in range(int(length)):
            d = self._draw_distance(seen, enable_padding, ra)
            if d == 0:
                line = stack.pop(0)
                seen += 1
            else:
                line = stack.pop(depth - d)
            stack.append(line)
            out[k] = line
        return out


def generate_synthetic_input_batch(m_den, ln_emb, n, num_indices_per_lookup, num_indices_per_lookup_fixed, trace_file,
                                   enable_padding=False, ra=np.random):
    """Like ``generate_uniform_input_batch`` with the ids of every bag replayed from table i's distribution file
    (``trace_file`` with its ``j`` replaced by i, as the reference does): bag size drawn first (unless fixed), then
    that many LRU-replayed ids, sorted and de-duplicated; ids outside the table are folded back by ``mod rows``."""
    X = torch.tensor(ra.rand(n, m_den).astype(np.float32))
    lS_o, lS_i = [], []
    for t, rows in enumerate(ln_emb):
        rows = int(rows)
        model = StackDistanceModel.read(trace_file.replace("j", str(t)))
        starts, ids, filled = np.empty(n, dtype=np.int64), [], 0
        for b in range(n):
            if num_indices_per_lookup_fixed:
                k = int(num_indices_per_lookup)
            else:
                u = ra.random_sample(1)
                k = int(max(1, np.round(u * min(rows, num_indices_per_lookup))[0]))
            bag = np.unique(model.generate_lru(k, enable_padding, ra)).astype(np.int64)
            if bag.min() < 0 or bag.max() >= rows:
                print("WARNING: distribution is inconsistent with embedding table size (using mod to recover and continue)")
                bag = np.mod(bag, rows).astype(np.int64)
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
        if data_generation not in ("random", "synthetic"):
            raise SystemExit("ERROR: --data-generation=" + data_generation + " is not supported")
        self.data_generation, self.trace_file, self.enable_padding = data_generation, trace_file, enable_padding
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
        if self.data_generation == "random":
            X, lS_o, lS_i = generate_uniform_input_batch(self.m_den, self.ln_emb, n, self.num_indices_per_lookup,
                                                         self.num_indices_per_lookup_fixed, ra=self.ra)
        else:
            X, lS_o, lS_i = generate_synthetic_input_batch(self.m_den, self.ln_emb, n, self.num_indices_per_lookup,
                                                           self.num_indices_per_lookup_fixed, self.trace_file,
                                                           self.enable_padding, ra=self.ra)
        T = generate_random_output_batch(n, self.num_targets, self.round_targets, ra=self.ra)
        return X, lS_o, lS_i, T


def collate_wrapper_random(list_of_tuples):
    """DataLoader(batch_size=1) collate: unwrap the single pre-built batch; offsets become one ``[T, B]`` tensor."""
    X, lS_o, lS_i, T = list_of_tuples[0]
    return X, torch.stack(lS_o), lS_i, T


def collate_wrapper_criteo(list_of_tuples):
    """DataLoader collate of the per-sample Criteo path (``T/dlrm_data_pytorch.py:339-352``): samples
    ``(X_int[13], X_cat[T], y)`` -> ``(log(X_int + 1) fp32 [B, 13], lS_o int64 [T, B] = arange(B) per table (one id per
    bag), lS_i int64 [T, B] = X_cat transposed, targets fp32 [B, 1])``.  One conversion per field instead of the
    reference's per-sample tensor construction; the whole-batch binary path does the same on the device
    (``cloudtik_b200.criteo_input``)."""
    x_int = np.asarray([t[0] for t in list_of_tuples])
    x_cat = np.asarray([t[1] for t in list_of_tuples])
    y = np.asarray([t[2] for t in list_of_tuples])
    X = torch.log(torch.as_tensor(x_int).to(torch.float) + 1)
    lS_i = torch.as_tensor(x_cat).to(torch.long).t().contiguous()
    B, T = x_cat.shape
    lS_o = torch.arange(B, dtype=torch.long).repeat(T, 1)
    return X, lS_o, lS_i, torch.as_tensor(y).to(torch.float32).view(-1, 1)


def make_random_data_and_loader(args, ln_emb, m_den, world_size=1):
    data = RandomDataset(m_den, ln_emb, args.data_size, args.num_batches, args.mini_batch_size,
                         args.num_indices_per_lookup, args.num_indices_per_lookup_fixed, 1, args.round_targets,
                         args.data_generation, getattr(args, "data_trace_file", ""),
                         getattr(args, "data_trace_enable_padding", False), reset_seed_on_access=True,
                         rand_seed=args.numpy_rand_seed, world_size=world_size)
    loader = torch.utils.data.DataLoader(data, batch_size=1, shuffle=False, num_workers=0,
                                         collate_fn=collate_wrapper_random, pin_memory=False, drop_last=False)
    return data, loader
