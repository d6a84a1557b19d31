"""Pin the CPU oracle (oracle/ops_numpy.py, oracle/dlrm_port.py) against the fixtures
written by oracle/gen_golden.py from the reference's own code (CPU-only tests)."""
import numpy as np
import torch

from oracle import ops_numpy as ops
from oracle.dlrm_port import DLRMPort, train_step


def test_interaction_kat(golden):
    g = golden("interaction")
    R = ops.interact_fwd(np.array([[1.0, 2.0]]), [np.array([[3.0, 4.0]]), np.array([[5.0, 6.0]])])
    np.testing.assert_array_equal(R, [[1, 2, 11, 17, 39]])
    np.testing.assert_array_equal(g["kat_R"], R)


def test_interaction_fwd_bwd(golden):
    g = golden("interaction")
    for tag in "abc":
        x, ly = g[f"{tag}_x"], list(g[f"{tag}_ly"])
        R = ops.interact_fwd(x, ly)
        np.testing.assert_allclose(R, g[f"{tag}_R"], rtol=1e-5, atol=1e-5)
        dx, dly = ops.interact_bwd(x, ly, g[f"{tag}_dR"])
        np.testing.assert_allclose(dx, g[f"{tag}_dx"], rtol=1e-5, atol=2e-5)
        np.testing.assert_allclose(np.stack(dly), g[f"{tag}_dly"], rtol=1e-5, atol=2e-5)


def test_index_kats(golden):
    g = golden("index_kats")
    idx, off = g["idx"], g["off"]
    np.testing.assert_array_equal(ops.offset2bag(off, len(idx)), g["offset2bag"])
    np.testing.assert_array_equal(ops.bag_lengths(off, len(idx)), g["bag_size"])
    skeys, perm, uniq, counts, seg = ops.sort_dedup(idx)
    np.testing.assert_array_equal(skeys, g["sorted"])
    np.testing.assert_array_equal(perm, g["perm"])
    np.testing.assert_array_equal(uniq, g["unique"])
    np.testing.assert_array_equal(uniq, g["coalesced_indices"])
    np.testing.assert_array_equal(counts, g["counts"])
    np.testing.assert_array_equal(seg, np.concatenate([[0], np.cumsum(g["counts"])]))
    ci, cv = ops.embedding_bag_sparse_grad(idx, off, g["gout"])
    np.testing.assert_array_equal(ci, g["coo_indices"])
    np.testing.assert_array_equal(cv, g["coo_values"])
    # coalesced values == per-unique-row sums
    acc = np.zeros((len(uniq), 4), np.float32)
    np.add.at(acc, np.searchsorted(uniq, ci), cv)
    np.testing.assert_allclose(acc, g["coalesced_values"], rtol=1e-6)


def test_partition(golden):
    for row in golden("index_kats")["partition"]:
        world, rank, n, start, stop, my_len = (int(v) for v in row[:6])
        sl = ops.get_my_slice(n, world, rank)
        assert (sl.start, sl.stop) == (start, stop)
        ml, splits = ops.get_split_lengths(n, world, rank)
        assert ml == my_len
        ref = [int(v) for v in row[6:6 + world]]
        assert (splits is None and ref[0] == -1) or splits == ref


def test_lr_policy(golden):
    g = golden("lr_policy")
    for c, (base, warm, start, nsteps) in enumerate(g["cfgs"]):
        last = None
        for step0, want in enumerate(g[f"trace{c}"]):
            # get_last_lr() after k scheduler steps was computed at _step_count = k+1
            lr = ops.lr_policy(step0 + 1, base, int(warm), int(start), int(nsteps), last)
            last = lr
            assert abs(lr - want) <= 1e-12 * max(1.0, abs(want)), (c, step0, lr, want)


def test_embedding_bag_and_sgd_match_torch():
    rs = np.random.RandomState(5)
    W = rs.randn(50, 8).astype(np.float32)
    idx = rs.randint(0, 50, size=37)
    off = np.array([0, 3, 3, 10, 10, 20, 37 - 4], dtype=np.int64)
    E = torch.nn.EmbeddingBag(50, 8, mode="sum", sparse=True, _weight=torch.tensor(W.copy()))
    out = E(torch.tensor(idx), torch.tensor(off))
    np.testing.assert_allclose(ops.embedding_bag_sum(W, idx, off), out.detach().numpy(), rtol=1e-6, atol=1e-6)
    gout = rs.randn(len(off), 8).astype(np.float32)
    out.backward(torch.tensor(gout))
    torch.optim.SGD(E.parameters(), lr=0.3).step()
    W2 = ops.sparse_sgd_update(W.copy(), idx, off, gout, 0.3)
    np.testing.assert_allclose(W2, E.weight.detach().numpy(), rtol=1e-6, atol=1e-6)


def test_adagrad_matches_torch():
    rs = np.random.RandomState(6)
    W = rs.randn(30, 4).astype(np.float32)
    E = torch.nn.EmbeddingBag(30, 4, mode="sum", sparse=True, _weight=torch.tensor(W.copy()))
    opt = torch.optim.Adagrad(E.parameters(), lr=0.01, eps=1e-10)
    state = np.zeros_like(W)
    Wn = W.copy()
    for _ in range(3):
        idx = rs.randint(0, 30, size=25)
        off = np.arange(0, 25, 5, dtype=np.int64)
        gout = rs.randn(5, 4).astype(np.float32)
        opt.zero_grad()
        E(torch.tensor(idx), torch.tensor(off)).backward(torch.tensor(gout))
        opt.step()
        ops.sparse_adagrad_update(Wn, state, idx, off, gout, 0.01, 1e-10)
    np.testing.assert_allclose(Wn, E.weight.detach().numpy(), rtol=1e-5, atol=1e-6)


def test_linearize():
    idx = [np.array([1, 2, 0]), np.array([4, 4]), np.array([], dtype=np.int64)]
    off = [np.array([0, 1]), np.array([0, 2]), np.array([0, 0])]
    i, o, w = ops.linearize_indices_and_offsets(idx, off, [3, 5, 7])
    np.testing.assert_array_equal(i, [1, 2, 0, 4, 4])
    np.testing.assert_array_equal(o, [0, 1, 3, 5, 5, 5, 5])
    np.testing.assert_array_equal(w, [1, 2, 0, 7, 7])


def _port_from_fixture(g):
    np.random.seed(int(g["seed"]))
    torch.manual_seed(int(g["seed"]))
    return DLRMPort(int(g["D"]), g["ln_emb"], g["ln_bot"], g["ln_top"])


def test_port_small_model_bit_exact(golden):
    g = golden("small_model")
    net = _port_from_fixture(g)
    for k, v in net.state_dict().items():
        np.testing.assert_array_equal(v.numpy(), g["init/" + k])
    opt = torch.optim.SGD(net.parameters(), lr=float(g["lr"]))
    nt = len(g["ln_emb"])
    losses = []
    for s in range(int(g["steps"])):
        lS_o = [torch.tensor(g[f"o{s}_{t}"]) for t in range(nt)]
        lS_i = [torch.tensor(g[f"i{s}_{t}"]) for t in range(nt)]
        losses.append(train_step(net, opt, torch.tensor(g[f"X{s}"]), lS_o, lS_i, torch.tensor(g[f"T{s}"])))
    np.testing.assert_array_equal(np.array(losses), g["losses"])
    for k, v in net.state_dict().items():
        np.testing.assert_array_equal(v.numpy(), g["final/" + k])


def test_port_kaggle_c1_losses(golden):
    """BASELINE config 1 shape: loss trace of the reference, reproduced by the port."""
    g = golden("kaggle_c1")
    net = _port_from_fixture(g)
    sums = np.array([net.emb_l[t].weight.detach().numpy().astype(np.float64).sum() for t in range(26)])
    np.testing.assert_array_equal(sums, g["init_table_sums"])
    opt = torch.optim.SGD(net.parameters(), lr=float(g["lr"]))
    losses = []
    for s in range(int(g["steps"])):
        lS_o = [torch.tensor(g["off"][s, t].astype(np.int64)) for t in range(26)]
        lS_i = [torch.tensor(g["idx"][s, t].astype(np.int64)) for t in range(26)]
        losses.append(train_step(net, opt, torch.tensor(g["X"][s]), lS_o, lS_i, torch.tensor(g["T"][s])))
    np.testing.assert_allclose(np.array(losses), g["losses"], rtol=1e-6)
    for t in (2, 23):
        np.testing.assert_allclose(net.emb_l[t].weight.detach().numpy()[g[f"rows_t{t}"]],
                                   g[f"final_rows_t{t}"], rtol=1e-6, atol=1e-8)
    np.testing.assert_allclose(net.top_l[0].weight.detach().numpy()[:8], g["final_top0_w"], rtol=1e-5, atol=1e-7)


def test_lr_scheduler_mirror_matches_reference_traces(golden):
    """cloudtik_b200.lr.LRPolicyScheduler (the product's host-side mirror) against the traces the
    reference's scheduler produced (tests/golden/lr_policy.npz)."""
    from cloudtik_b200.lr import LRPolicyScheduler
    g = golden("lr_policy")
    p = torch.nn.Parameter(torch.zeros(1))
    for c, (base, warm, start, nsteps) in enumerate(g["cfgs"]):
        opt = torch.optim.SGD([p], lr=float(base))
        sch = LRPolicyScheduler(opt, int(warm), int(start), int(nsteps))
        got = []
        for _ in range(20):
            got.append(sch.get_last_lr()[0])
            opt.step()
            sch.step()
        np.testing.assert_allclose(got, g[f"trace{c}"], rtol=1e-12)


def test_criteo_transform_matches_reference_fixture(golden):
    """oracle.criteo_transform vs T/data_loader_terabyte._transform_features on raw int32 records:
    ids / offsets / labels bit-exact (negative ids, ids beyond max_ind_range, max_ind_range = 0),
    dense = log(x + 1) within one float ulp of torch's log."""
    g = golden("criteo_transform")
    for tag in "abc":
        X, lS_o, lS_i, Y = ops.criteo_transform(g[f"{tag}_records"], int(g[f"{tag}_max_ind_range"]))
        np.testing.assert_array_equal(lS_i, g[f"{tag}_lS_i"])
        np.testing.assert_array_equal(lS_o, g[f"{tag}_lS_o"])
        np.testing.assert_array_equal(Y, g[f"{tag}_Y"])
        np.testing.assert_allclose(X, g[f"{tag}_X"], rtol=2e-7, atol=0)
        assert lS_i.min() >= 0 or int(g[f"{tag}_max_ind_range"]) == 0


def test_int8_rowwise_quantisation_matches_torch_observer(golden):
    """oracle.quantize_rowwise_int8 / embedding_bag_sum_int8 vs the fixture written with torch.ao's
    PerChannelMinMaxObserver + quantize_per_channel (the observers the reference's int8 QConfig names)."""
    g = golden("int8_tables")
    q, s = ops.quantize_rowwise_int8(g["W"])
    np.testing.assert_array_equal(q, g["q"])
    np.testing.assert_array_equal(s, g["scale"])
    assert not g["zero_point"].any()
    got = ops.embedding_bag_sum_int8(q, s, g["idx"], g["off"][:-1])
    np.testing.assert_allclose(got, g["pooled"], rtol=1e-6, atol=1e-6)


def test_md_solver_mirror_matches_reference(golden):
    """cloudtik_b200.tricks.md_solver (pure index arithmetic, CPU) vs T/tricks/md_embedding_bag.md_solver."""
    from cloudtik_b200.tricks import md_solver
    g = golden("tricks")
    n = torch.tensor(g["md_n"])
    for k, alpha in enumerate((0.0, 0.25, 0.5)):
        np.testing.assert_array_equal(md_solver(n.clone(), alpha, d0=128, round_dim=True).numpy(), g[f"md_d0_{k}"])
        np.testing.assert_array_equal(md_solver(n.clone(), alpha, d0=128, round_dim=False).numpy(), g[f"md_raw_{k}"])
    np.testing.assert_array_equal(md_solver(n.clone(), 0.3, B=1.0e8, round_dim=False).numpy(), g["md_budget"])


def test_product_side_quantiser_matches_fixture_on_cpu(golden):
    """cloudtik_b200.quantized.quantize_rowwise_int8 is plain torch arithmetic: on CPU tensors it must give the
    fixture's codes and scales bit for bit (the GPU run of the same function is checked in tests/test_gpu_inference.py)."""
    from cloudtik_b200.quantized import quantize_rowwise_int8
    g = golden("int8_tables")
    q, s = quantize_rowwise_int8(torch.tensor(g["W"]), chunk_rows=64)   # several chunks
    np.testing.assert_array_equal(q.numpy(), g["q"])
    np.testing.assert_array_equal(s.numpy(), g["scale"])


def _t_smoke_state(g, tag, which):
    return {str(n): torch.tensor(g[f"{tag}_{which}{k}"]) for k, n in enumerate(g[f"{tag}_param_names"])}


def _t_smoke_batches(g, tag):
    for s in range(int(g[f"{tag}_steps"])):
        yield (torch.tensor(g[f"{tag}_X{s}"]), [torch.tensor(g[f"{tag}_o{s}_{t}"]) for t in range(3)],
               [torch.tensor(g[f"{tag}_i{s}_{t}"]) for t in range(3)], torch.tensor(g[f"{tag}_T{s}"]))


def test_port_reproduces_the_reference_test_script_configs(golden):
    """The two configurations of the reference's own test script (T/test/dlrm_s_test.sh:24,26; MSE loss, ragged
    bags of up to 10 indices, D = 2): the fixture holds the losses the T/ model printed (0.028790, 0.156787, ...);
    the oracle port, started from the same weights, must print the same six decimals (losses within 1e-6)."""
    g = golden("t_smoke")
    assert [round(float(v), 6) for v in g["c2_losses"]] == [0.028790, 0.156787]
    assert [round(float(v), 6) for v in g["c4_losses"]] == [0.028790, 0.156787, 0.183348, 0.022413, 0.134285,
                                                           0.165428, 0.016494, 0.111174, 0.137763]
    for tag in ("c2", "c4"):
        np.random.seed(0)
        net = DLRMPort(2, g["ln_emb"], g["ln_bot"], g["ln_top"])
        net.load_state_dict(_t_smoke_state(g, tag, "init"))
        net.loss_fn = torch.nn.MSELoss(reduction="mean")
        opt = torch.optim.SGD(net.parameters(), lr=0.1)
        losses = [train_step(net, opt, *b) for b in _t_smoke_batches(g, tag)]
        # (the dense-vs-sparse summation order explained below reaches the losses from the second epoch on: 1e-8)
        np.testing.assert_allclose(np.array(losses), g[f"{tag}_losses"], rtol=1e-6, atol=0)
        assert [round(v, 6) for v in losses] == [round(float(v), 6) for v in g[f"{tag}_losses"]]
        # final weights: the T/ variant builds tables of < 2048 rows with sparse=False (T/dlrm_s_pytorch.py:266-269),
        # i.e. a dense gradient whose duplicate rows are summed in another order than the sparse path's: last-ulp
        for k, v in _t_smoke_state(g, tag, "final").items():
            np.testing.assert_allclose(net.state_dict()[k].numpy(), v.numpy(), rtol=1e-6, atol=1e-7)


def test_batch_producers_reproduce_the_reference_stream(golden):
    """cloudtik_b200.data (host-side mirror of T/dlrm_data_pytorch's random path): for the reference's seed and
    flags it must hand out the very batches the reference's RandomDataset + DataLoader produced (fixture t_smoke:
    ragged bags of up to 10 unique sorted ids, same batches again in every epoch), and the module-level generator
    must reproduce the small_model fixture drawn with the reference's generate_uniform_input_batch."""
    import types
    from cloudtik_b200 import data as D
    g = golden("t_smoke")
    for tag, data_size, nepochs in (("c2", 4, 1), ("c4", 5, 3)):
        args = types.SimpleNamespace(data_size=data_size, num_batches=0, mini_batch_size=2, num_indices_per_lookup=10,
                                     num_indices_per_lookup_fixed=False, round_targets=False, data_generation="random",
                                     numpy_rand_seed=123)
        _, ld = D.make_random_data_and_loader(args, g["ln_emb"], 4)
        s = 0
        for _ep in range(nepochs):
            for X, lS_o, lS_i, T in ld:
                np.testing.assert_array_equal(X.numpy(), g[f"{tag}_X{s}"])
                np.testing.assert_array_equal(T.numpy(), g[f"{tag}_T{s}"])
                for t in range(3):
                    np.testing.assert_array_equal(lS_o[t].numpy(), g[f"{tag}_o{s}_{t}"])
                    np.testing.assert_array_equal(lS_i[t].numpy(), g[f"{tag}_i{s}_{t}"])
                s += 1
        assert s == int(g[f"{tag}_steps"])
    sm = golden("small_model")
    ra = np.random.RandomState(int(sm["seed"]) + 1)
    for s in range(int(sm["steps"])):
        X, lS_o, lS_i = D.generate_uniform_input_batch(13, sm["ln_emb"], int(sm["B"]), 5, False, ra=ra)
        T = D.generate_random_output_batch(int(sm["B"]), 1, round_targets=True, ra=ra)
        np.testing.assert_array_equal(X.numpy(), sm[f"X{s}"])
        np.testing.assert_array_equal(T.numpy(), sm[f"T{s}"])
        for t in range(len(sm["ln_emb"])):
            np.testing.assert_array_equal(lS_o[t].numpy(), sm[f"o{s}_{t}"])
            np.testing.assert_array_equal(lS_i[t].numpy(), sm[f"i{s}_{t}"])


def test_trace_driven_producer_reproduces_the_reference(golden, tmp_path):
    """--data-generation=synthetic (cloudtik_b200.data.StackDistanceModel / generate_synthetic_input_batch) against
    tests/golden/trace_synth.npz, every array of which the reference's own trace_profile / trace_generate_lru /
    generate_synthetic_input_batch wrote (oracle/gen_golden.py gen_trace_synth): the profile of a trace (first-touch
    lines, distances, cumulative probabilities -- bit-equal floats), the LRU replay under the same seed, the file
    round trip, and whole batches (ragged, fixed bag size, padding on)."""
    import types
    from cloudtik_b200 import data as D
    g = golden("trace_synth")
    ln_emb = [int(x) for x in g["ln_emb"]]
    for t in range(len(ln_emb)):
        for pad in (False, True):
            tag = f"{t}_{int(pad)}"
            m = D.StackDistanceModel.from_trace(g[f"trace{t}"], pad)
            np.testing.assert_array_equal(np.array(m.lines), g[f"lines{tag}"])
            np.testing.assert_array_equal(np.array(m.distances), g[f"dist{tag}"])
            np.testing.assert_array_equal(m.cumulative, g[f"cumm{tag}"])
            path = str(tmp_path / f"dist{int(pad)}_emb_{t}.log")
            m.write(path)
            back = D.StackDistanceModel.read(path)
            assert back.lines == m.lines and back.distances == m.distances
            np.testing.assert_array_equal(back.cumulative, m.cumulative)  # repr() round-trips a double
            ids = back.generate_lru(300, pad, ra=np.random.RandomState(7 + t))
            np.testing.assert_array_equal(ids, g[f"lru{tag}"])
            assert m.lines == back.lines  # the replay works on a copy
    for tag, fixed, pad in (("ragged", False, False), ("fixed", True, False), ("ragged_pad", False, True)):
        X, lS_o, lS_i = D.generate_synthetic_input_batch(4, ln_emb, 6, 5, fixed, str(tmp_path / f"dist{int(pad)}_emb_j.log"),
                                                         pad, ra=np.random.RandomState(123))
        np.testing.assert_array_equal(X.numpy(), g[f"{tag}_X"])
        for t in range(len(ln_emb)):
            np.testing.assert_array_equal(lS_o[t].numpy(), g[f"{tag}_o{t}"])
            np.testing.assert_array_equal(lS_i[t].numpy(), g[f"{tag}_i{t}"])
    # through the dataset / loader the train CLI uses
    args = types.SimpleNamespace(data_size=12, num_batches=0, mini_batch_size=6, num_indices_per_lookup=5,
                                 num_indices_per_lookup_fixed=False, round_targets=False, data_generation="synthetic",
                                 data_trace_file=str(tmp_path / "dist0_emb_j.log"), data_trace_enable_padding=False,
                                 numpy_rand_seed=123)
    _, ld = D.make_random_data_and_loader(args, ln_emb, 4)
    X, lS_o, lS_i, T = next(iter(ld))
    np.testing.assert_array_equal(X.numpy(), g["ragged_X"])
    np.testing.assert_array_equal(lS_i[1].numpy(), g["ragged_i1"])
    assert T.shape == (6, 1)


def test_criteo_collate_matches_the_reference(golden):
    """cloudtik_b200.data.collate_wrapper_criteo against the batch the reference's function built from the same
    per-sample tuples (tests/golden/collate_criteo.npz): bit-equal dense features (log(x + 1) in fp32), offsets,
    transposed ids and targets, for list and for array samples."""
    from cloudtik_b200 import data as D
    g = golden("collate_criteo")
    for as_list in (True, False):
        samples = [((a.tolist(), b.tolist(), int(c)) if as_list else (a, b, c))
                   for a, b, c in zip(g["x_int"], g["x_cat"], g["y"])]
        X, lS_o, lS_i, T = D.collate_wrapper_criteo(samples)
        np.testing.assert_array_equal(X.numpy(), g["X"])
        np.testing.assert_array_equal(lS_o.numpy(), g["lS_o"])
        np.testing.assert_array_equal(lS_i.numpy(), g["lS_i"])
        np.testing.assert_array_equal(T.numpy(), g["T"])
        assert X.dtype == torch.float32 and lS_o.dtype == torch.int64 and lS_i.dtype == torch.int64
