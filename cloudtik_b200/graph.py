"""Whole-step CUDA graph: forward + loss + backward (incl. the fused sparse update) + dense SGD.

The step launches ~110 small kernels (torch MLP pieces) around the hand-written ones; replaying
them as one graph removes the launch gaps ("CUDA streams and graphs instead of a tracing
compiler").  Every C-ABI call is stream-ordered, allocation-free and sync-free, so it is captured
like any other kernel; the side-stream sort joins the capture through event fork/join.

Static input buffers are owned by this object: ``step(X, Y, indices, offsets)`` copies the batch
into them (device or pinned-host sources both work, non-blocking) and replays.
"""
from __future__ import annotations

import torch


class GraphedTrainStep:
    def __init__(self, net, optimizer, example, warmup: int = 3, post_backward=None, loss_scale: float = 1.0):
        """``example`` = (X, Y, indices, offsets) device tensors with the shapes of every later batch -- or, for a
        micro-batched step, (X, Y, indices_0, offsets_0, indices_1, offsets_1, ...), see ``dlrm.train_step_chunked``.
        ``post_backward``: optional callable run between backward and the optimizer step (the flat
        all-reduce of the MLP gradients in the multi-GPU case; NCCL collectives are captured too)."""
        self.net, self.opt = net, optimizer
        self.post_backward = post_backward
        self.loss_scale = loss_scale
        self.static = [torch.empty_like(t) for t in example]
        for d, s in zip(self.static, example):
            d.copy_(s)
        emb = net.emb_l
        if emb.profile is not None:
            raise RuntimeError("disable MergedEmbeddingBagWithSGD.profile before capturing a graph")
        # warm up on a side stream (allocator pools, cuBLAS handles / workspaces, the embedding
        # workspace, the tensor-map entry point) -- nothing may allocate lazily during capture
        # The step's main chain (MLPs, interaction) is captured on a HIGH-priority stream; the embedding's side
        # stream (index stage, HBM- / NVLink-bound exchange and update kernels) has the default, lowest priority.
        # Kernel nodes keep the priority of the stream they were captured on, so when both have thread blocks
        # waiting the block scheduler places the main chain's first: the side work fills the SM slots the main
        # chain leaves free instead of queueing thousands of blocks ahead of a small GEMM.
        s = torch.cuda.Stream(priority=-1)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(warmup):
                self._eager()
        torch.cuda.current_stream().wait_stream(s)
        emb.join()  # no event from outside the capture may be waited on inside it
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=s):
            self.loss = self._eager()
        torch.cuda.synchronize()

    def _eager(self):
        from .dlrm import train_step_chunked
        return train_step_chunked(self.net, self.opt, self.static, self.post_backward, self.loss_scale)

    # ---- host-fed operation: pinned-host batches are staged by a copy stream one step ahead ----
    def prefetch(self, host_batch):
        """Start the H2D copy of a pinned-host batch (X, Y, indices, offsets) into the staging
        buffers on the copy stream; ``step_prefetched()`` consumes it."""
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream()
            self._stage = [[torch.empty_like(t) for t in self.static] for _ in range(2)]
            self._stage_ready = [None, None]
            self._stage_free = [None, None]
            self._slot = 0
        k = self._slot
        with torch.cuda.stream(self._copy_stream):
            if self._stage_free[k] is not None:
                self._copy_stream.wait_event(self._stage_free[k])  # previous consumer of this slot is done
            for d, s in zip(self._stage[k], host_batch):
                d.copy_(s, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        self._stage_ready[k] = ev
        self._slot ^= 1
        return k

    def step_prefetched(self, k):
        """Replay on the batch staged in slot ``k`` (device-to-device hop into the static buffers)."""
        main = torch.cuda.current_stream()
        main.wait_event(self._stage_ready[k])
        for d, s in zip(self.static, self._stage[k]):
            d.copy_(s, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(main)
        self._stage_free[k] = ev
        self.graph.replay()
        return self.loss

    # ---- raw-record operation (cloudtik_b200.criteo_input): the host supplies int32 Criteo records ----
    def prefetch_records(self, host_records):
        """Start the H2D copy of a pinned-host batch of RAW records (int32 [B, tot_fea], 160 B per sample);
        ``step_records()`` decodes it on the device straight into the graph's static inputs."""
        if not hasattr(self, "_rec_stream"):
            self._rec_stream = torch.cuda.Stream()
            dev = self.static[0].device
            self._rec_stage = [torch.empty(host_records.shape, dtype=torch.int32, device=dev) for _ in range(2)]
            self._rec_ready = [None, None]
            self._rec_free = [None, None]
            self._rec_slot = 0
        k = self._rec_slot
        with torch.cuda.stream(self._rec_stream):
            if self._rec_free[k] is not None:
                self._rec_stream.wait_event(self._rec_free[k])
            self._rec_stage[k].copy_(host_records, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._rec_stream)
        self._rec_ready[k] = ev
        self._rec_slot ^= 1
        return k

    def step_records(self, k, max_ind_range: int = 0, **slices):
        """One ``b200_criteo_transform`` launch (records of slot ``k`` -> X, Y, indices, offsets in the static
        buffers: no device-to-device hop), then the graph replay.  ``slices``: tab_begin / tab_count /
        row_begin / row_count for a rank of a table-wise sharded job."""
        from . import ops
        main = torch.cuda.current_stream()
        main.wait_event(self._rec_ready[k])
        X, Y, idx, off = self.static[:4]
        ops.criteo_transform(self._rec_stage[k], max_ind_range, n_dense=X.shape[1], out=(X, Y, idx, off), **slices)
        ev = torch.cuda.Event()
        ev.record(main)
        self._rec_free[k] = ev
        self.graph.replay()
        return self.loss

    def loss_to_host_async(self):
        """Queue a device-to-host copy of the step's loss into a pinned slot and return a zero-argument reader.
        Calling the reader one step later costs nothing (the copy is long done) and keeps the GPU busy back to
        back -- the reference's per-step ``E.detach().cpu()`` (``P/dlrm_s_pytorch.py:1048``) stalls it every step."""
        if not hasattr(self, "_loss_host"):
            self._loss_host = [torch.empty((), dtype=self.loss.dtype).pin_memory() for _ in range(4)]
            self._loss_slot = 0
        k = self._loss_slot
        self._loss_slot = (k + 1) % len(self._loss_host)
        self._loss_host[k].copy_(self.loss, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()

        def read():
            ev.synchronize()
            return float(self._loss_host[k])
        return read

    def step(self, *batch):
        """Copy one batch (same tuple layout as ``example``) into the static buffers and replay; returns the
        (device) loss tensor."""
        for d, s in zip(self.static, batch):
            if d.data_ptr() != s.data_ptr():
                d.copy_(s, non_blocking=True)
        self.graph.replay()
        return self.loss
