"""Executable specification of the update kernel's run bookkeeping (cloudtik_b200/csrc/emb_bwd.cu), on the CPU.

The CUDA kernel cuts the SORTED positions into windows of C, NG windows per CTA, and decides with a handful of
key comparisons which group finishes which run: inside a window, through the in-block join (pieces in shared
memory, slot 0 / slot 1), or through the per-CTA partial rows that window_join_kernel adds.  This model follows
those rules literally with scalar "gradients" and must reproduce plain segment sums for ANY run structure --
a property test of the algorithm, independent of the GPU (the GPU tests check the CUDA code against the oracle).
"""
import numpy as np
from hypothesis import given, settings, strategies as st


def model_update(keys, vals, C, NG):
    """keys: sorted int array; vals: one number per position.  Returns {key: sum} computed the kernel's way, and
    checks on the way that every run is finished exactly once."""
    n = len(keys)
    done = {}
    BC = C * NG
    nblk = -(-n // BC)
    partial = {}  # (block, slot) -> value

    def finish_row(key, value):
        assert key not in done, "row written twice"
        done[key] = value

    for blk in range(nblk):
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

    # window_join_kernel: one group per CTA border; the border a run crosses FIRST owns it
    for kb in range(1, nblk):
        if keys[kb * BC - 1] != keys[kb * BC]:
            continue
        key = keys[kb * BC]
        s = int(np.searchsorted(keys, key, side="left"))
        e = int(np.searchsorted(keys, key, side="right"))
        if s // BC + 1 != kb:
            continue
        acc = 0.0
        for i in range(s // BC, (e - 1) // BC + 1):
            acc += partial.pop((i, 1 if s > i * BC else 0))
        finish_row(key, acc)
    assert not partial, "a per-CTA partial row was never consumed"
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
