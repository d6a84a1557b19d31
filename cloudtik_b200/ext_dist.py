"""Table-wise model-parallel plumbing: one process per GPU over torch.distributed (NCCL).

Mirror of the reference's ``extend_distributed`` module (``P/extend_distributed.py``): same
public names and argument meaning -- ``init_distributed``, ``get_my_slice``,
``get_split_lengths``, ``alltoall(inputs, per_rank_split_lengths) -> Request`` with
``Request.wait()``, ``all_gather``, ``barrier``, ``DDP`` -- so the model code reads like the
reference's.  Differences, on purpose:

* rank/size come from ``torch.distributed`` (the reference's module-global ``my_rank`` is never
  set: quirk q1 in SURVEY.md section 7.1);
* one ``all_to_all_single`` implementation (NCCL, or gloo in the CPU tests); the scatter /
  scatter-list fall-backs for back-ends without all-to-all (``P/extend_distributed.py:132-269``)
  are not needed;
* the forward all-to-all runs on a side stream so that it overlaps the bottom MLP, and the
  send buffer is the pooled-embedding buffer itself when the embedding kernel wrote it in the
  ``[N, lS*E]`` send layout (no ``torch.cat`` copy, ``P/extend_distributed.py:280``).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
from torch.autograd import Function
from torch.nn.parallel import DistributedDataParallel as DDP  # noqa: F401  (re-exported)

my_rank = -1
my_size = -1
my_local_rank = -1
my_local_size = -1
alltoall_supported = False
myreq = None


def env2int(env_list, default=-1):
    # P/extend_distributed.py:27-31
    for e in env_list:
        val = int(os.environ.get(e, -1))
        if val >= 0:
            return val
    return default


def _world():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def _rank():
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def get_my_slice(n):
    # P/extend_distributed.py:33-37
    size, rank = _world(), _rank()
    k, m = divmod(n, size)
    return slice(rank * k + min(rank, m), (rank + 1) * k + min(rank + 1, m), 1)


def get_split_lengths(n):
    # P/extend_distributed.py:39-49
    size = _world()
    k, m = divmod(n, size)
    if m == 0:
        return k, None
    splits = [(k + 1) if i < m else k for i in range(size)]
    return splits[_rank()], splits


def init_distributed(rank=-1, size=-1, backend=""):
    """P/extend_distributed.py:51-118, reduced to the launchers this build supports
    (torchrun / cloudtik-run env: RANK, WORLD_SIZE, LOCAL_RANK, MASTER_ADDR, MASTER_PORT)."""
    global myreq, my_rank, my_size, my_local_rank, my_local_size, alltoall_supported
    if size == -1:
        size = env2int(["PMI_SIZE", "OMPI_COMM_WORLD_SIZE", "MV2_COMM_WORLD_SIZE", "WORLD_SIZE"], 1)
    if rank == -1:
        rank = env2int(["PMI_RANK", "OMPI_COMM_WORLD_RANK", "MV2_COMM_WORLD_RANK", "RANK"], 0)
    if backend == "":
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        os.environ.setdefault("RANK", str(rank))
        os.environ.setdefault("WORLD_SIZE", str(size))
        my_local_rank = env2int(["MPI_LOCALRANKID", "OMPI_COMM_WORLD_LOCAL_RANK", "MV2_COMM_WORLD_LOCAL_RANK",
                                 "LOCAL_RANK"], 0)
        my_local_size = env2int(["MPI_LOCALNRANKS", "OMPI_COMM_WORLD_LOCAL_SIZE", "MV2_COMM_WORLD_LOCAL_SIZE",
                                 "LOCAL_WORLD_SIZE"], 1)
        if backend == "nccl":
            torch.cuda.set_device(my_local_rank)
        if not dist.is_initialized():
            dist.init_process_group(backend, rank=rank, world_size=size)
        my_rank = dist.get_rank()
        my_size = dist.get_world_size()
        alltoall_supported = True
        if my_rank == 0:
            print("Running on %d ranks using %s backend" % (my_size, backend))
    else:
        my_rank, my_size, my_local_rank, my_local_size = 0, 1, 0, 1
    myreq = Request()


class All2AllInfo(object):
    pass


class Request(object):
    """One outstanding all-to-all (the reference keeps a module-global one too,
    P/extend_distributed.py:25,120-130): not re-entrant, one per process."""

    def __init__(self):
        self.req = None
        self.tensor = None
        self.a2ai = None
        self.event = None

    def wait(self):
        ret = All2All_Wait.apply(*self.tensor)
        self.req = None
        self.tensor = None
        return ret


def _finish(req_holder):
    if req_holder.req is not None:
        req_holder.req.wait()
        req_holder.req = None


def _send_view(inputs, N, E):
    """[N, lS*E] send buffer: the tensors' common base if they are adjacent [N, E] column
    blocks of one row-major buffer (what the merged embedding kernel writes), else a cat."""
    t0 = inputs[0]
    lS = len(inputs)
    st = t0.untyped_storage().data_ptr()
    ok = all(t.untyped_storage().data_ptr() == st and tuple(t.shape) == (N, E) and t.stride() == (lS * E, 1)
             and t.storage_offset() == t0.storage_offset() + k * E for k, t in enumerate(inputs))
    if ok:
        return torch.as_strided(t0, (N * lS * E,), (1,), t0.storage_offset())
    return torch.cat(inputs, dim=1).view([-1])


class All2All_Req(Function):
    # P/extend_distributed.py:271-305
    @staticmethod
    def forward(ctx, a2ai, *inputs):
        global myreq
        mb_split_lengths = a2ai.gNS
        if mb_split_lengths:
            mb_split_lengths = [m * a2ai.lS * a2ai.E for m in mb_split_lengths]
        emb_split_lengths = a2ai.gSS
        if emb_split_lengths:
            emb_split_lengths = [a2ai.lN * e * a2ai.E for e in emb_split_lengths]
        input = _send_view(inputs, a2ai.N, a2ai.E)
        output = input.new_empty([a2ai.S * a2ai.lN * a2ai.E])
        myreq.req = dist.all_to_all_single(output, input, emb_split_lengths, mb_split_lengths, async_op=True)
        myreq.tensor = (output,)
        a2ai.mb_split_lengths = mb_split_lengths
        a2ai.emb_split_lengths = emb_split_lengths
        myreq.a2ai = a2ai
        ctx.a2ai = a2ai
        return myreq.tensor

    @staticmethod
    def backward(ctx, *grad_output):
        global myreq
        a2ai = ctx.a2ai
        _finish(myreq)
        grad_input = myreq.tensor
        myreq.tensor = None
        # [N, lS*E] -> per-table [N, E] views (no copy: the fused embedding backward reads them
        # in place through their common base)
        gi = grad_input.view([a2ai.N, a2ai.lS, a2ai.E])
        return (None, *[gi[:, k] for k in range(a2ai.lS)])


class All2All_Wait(Function):
    # P/extend_distributed.py:307-333
    @staticmethod
    def forward(ctx, *output):
        global myreq
        a2ai = myreq.a2ai
        ctx.a2ai = a2ai
        _finish(myreq)
        myreq.tensor = None
        emb_split_lengths = a2ai.emb_split_lengths if a2ai.emb_split_lengths else a2ai.lS * a2ai.lN * a2ai.E
        outputs = output[0].split(emb_split_lengths)
        return tuple(out.view([a2ai.lN, -1]) for out in outputs)

    @staticmethod
    def backward(ctx, *grad_outputs):
        global myreq
        a2ai = ctx.a2ai
        grad_output = torch.cat([g.contiguous().view([-1]) for g in grad_outputs])
        grad_input = grad_output.new_empty([a2ai.N * a2ai.lS * a2ai.E])
        myreq.req = dist.all_to_all_single(grad_input, grad_output, a2ai.mb_split_lengths,
                                           a2ai.emb_split_lengths, async_op=True)
        myreq.tensor = grad_input
        return (grad_output,)


def alltoall(inputs, per_rank_split_lengths):
    """P/extend_distributed.py:386-409.  ``inputs``: this rank's lS pooled tensors [N, E] for the
    GLOBAL batch; returns the Request whose ``wait()`` yields one [lN, lS_r*E] tensor per rank."""
    global myreq
    N, E = inputs[0].size()
    a2ai = All2AllInfo()
    a2ai.lS = len(inputs)
    a2ai.gSS = per_rank_split_lengths
    a2ai.lN, a2ai.gNS = get_split_lengths(N)
    a2ai.E = E
    a2ai.N = N
    a2ai.S = sum(per_rank_split_lengths) if per_rank_split_lengths else a2ai.lS * my_size
    All2All_Req.apply(a2ai, *inputs)
    return myreq


def shuffle_data(inputs):
    """T/extend_distributed.py:408-413 (call site T/dlrm_s_pytorch.py:494-497): every rank holds the
    indices of ALL tables for ITS local batch (a list of T tensors [B_loc]); one equal-split all-to-all
    hands rank r the indices of ITS tables for every rank's batch.  Returns ``[world, lS*B_loc]``: row s =
    rank s's local batch for my lS = T/world tables, so ``out[:, i*B_loc:(i+1)*B_loc].reshape(-1)`` is table
    i's index list for the GLOBAL batch (one-hot only, like the reference: offsets are an arange)."""
    input = torch.cat(inputs)
    if input.numel() % my_size != 0:
        raise ValueError("shuffle_data: the number of tables must be a multiple of the number of ranks")
    output = input.new_empty(input.size())
    dist.all_to_all_single(output, input)
    return output.reshape(my_size, -1)


def shuffle_bags(lS_o, lS_i, per_rank_tables=None, fixed_pooling=None):
    """Index all-to-all of the training variant (T/extend_distributed.py:408-413, call site
    T/dlrm_s_pytorch.py:494-497), generalised in the two ways SURVEY.md quirk q2 asks for: the number of tables
    need not be a multiple of the number of ranks (``per_rank_tables`` = ``get_split_lengths(T)[1]``, None = equal),
    and bags may hold any number of indices (the bag LENGTHS travel too; the reference regenerates the offsets as
    an arange, i.e. supports one index per bag only).

    Every rank passes the bags of ALL ``T`` sharded tables for ITS local batch: ``lS_o[t]`` int64[B_loc] start
    offsets, ``lS_i[t]`` int64[nnz_t].  Returns ``(g_o, g_i)`` for the tables THIS rank owns (contiguous partition,
    ``get_my_slice``): bags of the GLOBAL batch in rank-major order -- what ``distributed_forward`` feeds its local
    tables.  ``fixed_pooling=P`` (every bag of every rank holds exactly P indices) needs no host synchronisation;
    otherwise the receive sizes are read back once."""
    world, rank = dist.get_world_size(), dist.get_rank()
    T, B, dev = len(lS_i), lS_o[0].numel(), lS_i[0].device
    counts = list(per_rank_tables) if per_rank_tables else None
    if counts is None:
        if T % world:
            raise ValueError("shuffle_bags: pass per_rank_tables when the tables do not divide evenly")
        counts = [T // world] * world
    first = [sum(counts[:r]) for r in range(world)]
    mine = counts[rank]
    nnz = [int(i.numel()) for i in lS_i]
    lens = torch.stack([torch.cat([o[1:], o.new_tensor([n])]) - o for o, n in zip(lS_o, nnz)]).to(torch.int64)  # [T, B]
    send_len = lens.reshape(-1)  # table-major = destination-major (each rank owns a contiguous run of tables)
    recv_len = send_len.new_empty(world * mine * B)
    dist.all_to_all_single(recv_len, send_len, [mine * B] * world, [c * B for c in counts])
    recv_len = recv_len.view(world, mine, B)
    send_idx = torch.cat([i.to(torch.int64) for i in lS_i]) if T else torch.empty(0, dtype=torch.int64, device=dev)
    in_splits = [sum(nnz[first[r]:first[r] + counts[r]]) for r in range(world)]
    if fixed_pooling is not None:
        P = int(fixed_pooling)
        blocks = [[B * P] * mine for _ in range(world)]
    else:
        blocks = recv_len.sum(dim=2).tolist()  # [world][mine] indices per (source rank, local table): one host sync
    out_splits = [int(sum(b)) for b in blocks]
    recv_idx = send_idx.new_empty(sum(out_splits))
    dist.all_to_all_single(recv_idx, send_idx, out_splits, in_splits)
    g_o, g_i = [], []
    if fixed_pooling is not None:
        v = recv_idx.view(world, mine, B * P)
        off = torch.arange(world * B, dtype=torch.int64, device=dev) * P
        for i in range(mine):
            g_i.append(v[:, i].reshape(-1))
            g_o.append(off)
        return g_o, g_i
    src_start = [sum(out_splits[:s]) for s in range(world)]
    for i in range(mine):
        parts = []
        for s in range(world):
            a = src_start[s] + int(sum(blocks[s][:i]))
            parts.append(recv_idx[a:a + int(blocks[s][i])])
        g_i.append(torch.cat(parts))
        li = recv_len[:, i].reshape(-1)
        g_o.append(torch.cumsum(li, 0) - li)
    return g_o, g_i


def global_batch_indices(shuffled, n_local_tables):
    """Per-table GLOBAL-batch index lists + the shared arange offsets from ``shuffle_data``'s output
    (T/dlrm_s_pytorch.py:495-497)."""
    b_loc = shuffled.shape[1] // n_local_tables
    g_i = [shuffled[:, i * b_loc:(i + 1) * b_loc].reshape(-1) for i in range(n_local_tables)]
    offset = torch.arange(b_loc * my_size, device=shuffled.device)
    return [offset for _ in range(n_local_tables)], g_i


class AllGather(Function):
    # P/extend_distributed.py:335-381 (dim 0 only: the eval path gathers [B_local, 1] scores)
    @staticmethod
    def forward(ctx, input, global_lengths):
        rank = dist.get_rank()
        ctx.start = sum(global_lengths[:rank])
        ctx.length = global_lengths[rank]
        input = input.contiguous()
        size = list(input.size())
        size[0] = sum(global_lengths)
        output = input.new_empty(size)
        dist.all_gather(list(output.split(global_lengths, dim=0)), input)
        return output

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.narrow(0, ctx.start, ctx.length), None


def all_gather(input, lengths, dim=0):
    assert dim == 0
    if not lengths:
        lengths = [input.size(0)] * my_size
    return AllGather.apply(input, lengths)


def barrier():
    if my_size > 1:
        dist.barrier()


class DenseGradAllreduce(object):
    """Data-parallel gradient averaging for the (small, 2.4 M-parameter) MLPs without wrapping them in
    DDP: one flat fp32 buffer, ONE all-reduce after backward.  Same result as the reference's
    ``ext_dist.DDP(bot_l) / DDP(top_l)`` (``P/dlrm_s_pytorch.py:955-957``: mean over ranks); used when the
    MLPs run through the fused bf16 path, whose autograd functions bypass ``DDP.forward``."""

    def __init__(self, params, average: bool = True, early=None, symmetric_tail: bool = False):
        """``average=False`` sums instead (use it when every rank already divides its loss by the world
        size -- the way to keep a scale-sensitive table optimiser such as Adagrad exact, see tests/dist_parity.py).

        ``early``: parameters whose gradients are complete well before the end of backward (the top MLP: 93 % of
        the dense parameters of a DLRM).  Their part of the buffer is all-reduced asynchronously as soon as the
        last of them has its gradient (post-accumulate hooks), i.e. under the interaction backward, the embedding
        update and the bottom-MLP backward, the way DDP overlaps its buckets; ``__call__`` then only reduces the
        rest and waits.  Call ``begin(n)`` before a step whose backward runs in n passes (micro-batches).

        ``symmetric_tail`` (one NVLink box, NCCL backend): that end-of-step remainder (the bottom MLP and replicated
        small tables: ~2 MB) lives in a symmetric-memory buffer and is reduced by torch's two-shot peer-memory
        all-reduce (reduce-scatter + all-gather through plain peer loads / stores, two signal-pad barriers) instead of
        an NCCL call whose 8-rank latency (~60 us) sits on the critical path at the very end of the step."""
        self.average = average
        early = [p for p in (early or []) if p.requires_grad]
        early_ids = {id(p) for p in early}
        rest = [p for p in params if p.requires_grad and id(p) not in early_ids]
        self.params = early + rest
        self.n_early = len(early)
        dev = self.params[0].device
        n_early = sum(p.numel() for p in early)
        n_rest = sum(p.numel() for p in rest)
        self.early_numel = n_early
        self._symm_group = None
        self.flat_rest = None
        if symmetric_tail and my_size > 1 and n_rest > 0 and dist.get_backend() == "nccl":
            import torch.distributed._symmetric_memory as symm
            pad = -(-n_rest // (1024 * my_size)) * (1024 * my_size)  # whole 16-byte vectors per rank's share
            self.flat_rest = symm.empty(pad, dtype=torch.float32, device=dev)
            self._symm_group = dist.group.WORLD.group_name
            symm.rendezvous(self.flat_rest, self._symm_group)
            self.flat_rest.zero_()
            self.flat = torch.zeros(max(n_early, 1), dtype=torch.float32, device=dev)
        else:
            self.flat = torch.zeros(n_early + n_rest, dtype=torch.float32, device=dev)
        self.views = []
        o = 0
        for k, p in enumerate(self.params):
            if k == self.n_early:
                o = 0 if self.flat_rest is not None else n_early
            buf = self.flat if (k < self.n_early or self.flat_rest is None) else self.flat_rest
            self.views.append(buf[o:o + p.numel()].view_as(p))
            o += p.numel()
        self._pending = 0
        self._passes = 1
        self._work = None
        self._avg_op = None
        if self.n_early and my_size > 1:
            for p in early:
                p.register_post_accumulate_grad_hook(self._on_grad)
            self.begin(1)

    def _op(self):
        if self._avg_op is None:
            nccl = dist.get_backend() == "nccl"
            self._avg_op = dist.ReduceOp.AVG if (self.average and nccl) else dist.ReduceOp.SUM
        return self._avg_op

    def begin(self, passes: int = 1):
        self._passes = max(1, int(passes))
        self._pending = self.n_early * self._passes

    def _on_grad(self, _p):
        if self._pending <= 0:
            return
        self._pending -= 1
        if self._pending == 0:
            k = self.n_early
            torch._foreach_copy_(self.views[:k], [p.grad for p in self.params[:k]])
            self._work = dist.all_reduce(self.flat[:self.early_numel], op=self._op(), async_op=True)

    def __call__(self):
        if my_size <= 1:
            return
        k = self.n_early if self._work is not None else 0
        rest = self.params[k:]
        if rest:
            grads = [p.grad if p.grad is not None else torch.zeros_like(p) for p in rest]
            torch._foreach_copy_(self.views[k:], grads)
            if self.flat_rest is not None:
                if k == 0 and self.n_early:  # the early bucket was not reduced from inside backward this time
                    dist.all_reduce(self.flat[:self.early_numel], op=self._op())
                torch.ops.symm_mem.two_shot_all_reduce_(self.flat_rest, "sum", self._symm_group)
                if self.average:
                    self.flat_rest.mul_(1.0 / my_size)
            else:
                dist.all_reduce(self.flat[self.early_numel if k else 0:], op=self._op())
        if self._work is not None:
            self._work.wait()
            self._work = None
        if self.average and self._op() != dist.ReduceOp.AVG:
            self.flat.mul_(1.0 / my_size)
        for p, v in zip(self.params, self.views):
            p.grad = v  # the optimizer reads the averaged gradient straight from the flat buffer
        self.begin(self._passes)
