"""CPU oracle for the DLRM training hot path -- TEST INFRASTRUCTURE, NOT PRODUCT.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import anything from this package, and only as the
checker (or as the timed CPU baseline).  The product path (``cloudtik_b200``)
never imports it and has no CPU fallback.

Contents
--------
``ops_numpy``   numpy restatement of each operator on the path (sum-pooled
                EmbeddingBag, offsets->bag bucketing, stable sort / dedup of row
                ids, sparse SGD / Adagrad row update, pairwise-dot interaction
                forward and backward, table partitioning helpers).  Every
                function cites the reference file:line it follows.
``dlrm_port``   torch-CPU restatement of the reference ``DLRM_Net`` training
                step (same ATen operators the reference calls) -- this is the
                model-level oracle and the timed CPU baseline.
``gen_golden``  script that imports the *real* reference from ``/root/reference``
                (build container only) and writes the known-answer fixtures
                under ``tests/golden/``.

Parity status: the reference stores no golden vectors for this path (its only
numerics test diffs PyTorch against Caffe2 at run time), so the
oracle is pinned against outputs of the reference itself, generated in the
build container by ``gen_golden.py`` and committed under ``tests/golden``.
The arithmetic itself lives in PyTorch ATen (reference pins torch==1.13.1,
this image has torch 2.11.0: same operator semantics).
"""
