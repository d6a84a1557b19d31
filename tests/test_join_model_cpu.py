"""Executable specification of the update kernel's run bookkeeping (cloudtik_b200/csrc/emb_bwd.cu), on the CPU.

The CUDA kernel cuts the SORTED positions into windows of C, NG windows per CTA, and decides with a handful of
key comparisons which group finishes which run: inside a window, through the in-block join (pieces in shared
memory, slot 0 / slot 1), or through the per-chunk partial rows and the ticket protocol of the cross-block join
(every chunk a run touches draws a ticket on the counter of the run's head chunk; the last drawer adds the partial
rows in chunk order).  Chunks are worked on by blocks in ANY order -- one chunk per block, or a wave of blocks that
claim chunks from a counter and take the tickets of up to ``kflush`` chunks at a time (the persistent launch).  This
model follows those rules literally with scalar "gradients" and must reproduce plain segment sums for ANY run
structure and ANY schedule -- a property test of the algorithm, independent of the GPU (the GPU tests check the CUDA
code against the oracle).
"""
import numpy as np
from hypothesis import given, settings, strategies as st


def model_update(keys, vals, C, NG, n_ctas=None, kflush=1, rs=None):
    """keys: sorted int array; vals: one number per position.  Returns {key: sum} computed the kernel's way, and
    checks on the way that every run is finished exactly once.  ``n_ctas``: size of the persistent wave (None: one
    chunk per block, in index order); ``rs``: RandomState that interleaves the blocks' actions."""
    n = len(keys)
    done = {}
    BC = C * NG
    nblk = -(-n // BC)
    partial = {}  # (block, slot) -> value
    records = {}  # chunk -> (entering run or None, leaving run or None), a run = (key, start, end)

    def finish_row(key, value):
        assert key not in done, "row written twice"
        done[key] = value

    def run_of(pos):
        key = keys[pos]
        return (key, int(np.searchsorted(keys, key, side="left")), int(np.searchsorted(keys, key, side="right")))

    def process_chunk(blk):
        first = blk * BC
        ngv = min(NG, -(-(n - first) // C))
        piece = {}   # (group, slot) -> value
        info = []    # per group: cnt, cont0, open_r, tail_slot, keys of the window
        for g in range(ngv):
            a = first + g * C
            cnt = min(C, n - a)
            k = keys[a:a + cnt]
            before = keys[a - 1] if a > 0 else None
            after = keys[a + cnt] if a + cnt < n else None
            cont0 = before is not None and before == k[0]
            cont, cur_head, acc = cont0, (-1 if cont0 else 0), 0.0
            for pos in range(cnt):
                if pos > 0 and k[pos - 1] != k[pos]:
                    cont, cur_head = False, pos
                acc += vals[a + pos]
                nxt = k[pos + 1] if pos + 1 < cnt else after
                if nxt is None or nxt != k[pos]:          # the run ends at this position
                    if not cont:
                        finish_row(k[pos], acc)            # began and ended inside the window
                    else:
                        piece[(g, 0)] = acc                # continuation that ends here
                    acc = 0.0
            open_r = after is not None and after == k[cnt - 1]
            tail_slot = 0 if cur_head <= 0 else 1
            if open_r:
                piece[(g, tail_slot)] = acc
            info.append((cnt, cont0, open_r, tail_slot, k))
        last_cnt, _, _, _, last_k = info[ngv - 1]
        a_last = first + (ngv - 1) * C + last_cnt
        key_after_block = keys[a_last] if a_last < n else None

        def finish(g, slot, key, from_left, is_open):
            acc = piece[(g, slot)]
            j = g + 1
            if is_open:
                while j < ngv and info[j][4][0] == key:
                    acc += piece[(j, 0)]
                    j += 1
            to_right = is_open and j == ngv and key_after_block is not None and key_after_block == key
            if from_left or to_right:
                bslot = 0 if (from_left or (g == 0 and slot == 0)) else 1
                assert (blk, bslot) not in partial
                partial[(blk, bslot)] = acc
            else:
                finish_row(key, acc)

        for g in range(ngv):
            cnt, cont0, open_r, tail_slot, k = info[g]
            pass_through = cont0 and open_r and k[0] == k[cnt - 1]
            if g == 0 and cont0:
                finish(0, 0, k[0], True, pass_through)
            if open_r and not pass_through:
                finish(g, tail_slot, k[cnt - 1], False, True)

        x_enters = first > 0 and keys[first - 1] == keys[first]
        x_leaves = key_after_block is not None and key_after_block == last_k[last_cnt - 1]
        if x_enters or x_leaves:
            records[blk] = (run_of(first) if x_enters else None, run_of(a_last - 1) if x_leaves else None)
        return x_enters or x_leaves

    # ---- cross-block join: tickets on the head chunk's counter, the last drawer adds the partial rows in chunk order
    blk_cnt = {}

    def ticket(run):
        key, s0, e0 = run
        h, w1 = s0 // BC, (e0 - 1) // BC
        blk_cnt[h] = blk_cnt.get(h, 0) + 1
        if blk_cnt[h] == w1 - h + 1:
            blk_cnt[h] = 0
            acc = 0.0
            for i in range(h, w1 + 1):  # every contributor has written its row before it drew: all are there
                acc += partial.pop((i, 1 if s0 > i * BC else 0))
            finish_row(key, acc)

    def flush(chunks):
        for blk in chunks:
            ent, lea = records.pop(blk)
            if ent is not None:
                ticket(ent)
            if lea is not None and (ent is None or lea != ent):
                ticket(lea)

    if n_ctas is None:
        for blk in range(nblk):
            if process_chunk(blk):
                flush([blk])
    else:
        rs = rs or np.random.RandomState(0)
        grid = min(n_ctas, max(nblk, 1))
        nxt = [grid]  # the claim counter hands out the chunks behind the first wave
        state = [{"blk": c, "pend": []} for c in range(grid)]
        live = list(range(grid))
        while live:
            c = live[rs.randint(len(live))]
            st_ = state[c]
            if st_["blk"] is not None and st_["blk"] < nblk:
                if process_chunk(st_["blk"]):
                    st_["pend"].append(st_["blk"])
                st_["blk"] = nxt[0]
                nxt[0] += 1
                if len(st_["pend"]) >= kflush:
                    flush(st_["pend"])
                    st_["pend"] = []
            else:
                flush(st_["pend"])
                live.remove(c)
    assert not partial, "a per-chunk partial row was never consumed"
    assert not records and all(v == 0 for v in blk_cnt.values()), "tickets left behind"
    return done


@settings(max_examples=300, deadline=None)
@given(runs=st.lists(st.integers(min_value=1, max_value=70), min_size=1, max_size=25),
       C=st.sampled_from([2, 3, 4, 8]), NG=st.sampled_from([1, 2, 3, 4]), seed=st.integers(0, 2**31 - 1))
def test_every_run_is_finished_once_with_the_right_sum(runs, C, NG, seed):
    rs = np.random.RandomState(seed)
    keys = np.concatenate([np.full(r, 10 * i + 7) for i, r in enumerate(runs)])
    vals = rs.randint(-5, 6, size=len(keys)).astype(np.float64)
    got = model_update(keys, vals, C, NG)
    want = {}
    for k, v in zip(keys.tolist(), vals.tolist()):
        want[k] = want.get(k, 0.0) + v
    assert got == want


def test_model_on_the_shapes_the_gpu_test_uses():
    lens = [31, 1, 32, 33, 255, 1, 256, 257, 3, 511, 513, 1, 1, 1024, 7, 2047, 2, 2049, 700, 5000, 9, 64, 64, 64, 1, 8191]
    keys = np.concatenate([np.full(n, r) for r, n in enumerate(lens)])
    vals = np.ones(len(keys))
    got = model_update(keys, vals, 32, 8)
    assert got == {r: float(n) for r, n in enumerate(lens)}


@settings(max_examples=300, deadline=None)
@given(runs=st.lists(st.integers(min_value=1, max_value=90), min_size=1, max_size=30),
       C=st.sampled_from([2, 3, 4]), NG=st.sampled_from([1, 2, 4]), n_ctas=st.integers(1, 7),
       kflush=st.integers(1, 5), seed=st.integers(0, 2**31 - 1))
def test_any_schedule_of_the_persistent_launch_gives_the_same_sums(runs, C, NG, n_ctas, kflush, seed):
    """Blocks claim chunks in any interleaving and take their tickets late (up to kflush chunks at a time): every run
    is still finished exactly once, by whoever draws its last ticket, and every counter is back at zero."""
    rs = np.random.RandomState(seed)
    keys = np.concatenate([np.full(r, 10 * i + 7) for i, r in enumerate(runs)])
    vals = rs.randint(-5, 6, size=len(keys)).astype(np.float64)
    got = model_update(keys, vals, C, NG, n_ctas=n_ctas, kflush=kflush, rs=rs)
    want = {}
    for k, v in zip(keys.tolist(), vals.tolist()):
        want[k] = want.get(k, 0.0) + v
    assert got == want
