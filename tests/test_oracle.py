"""The oracle against closed-form identities and hypothesis properties (its parity with
TensorFlow 1.3 itself is unpinned -- see oracle/oracle.c)."""
import numpy as np
import pytest
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import oracle as O


def test_dense_apply_matches_numpy_float32_step_by_step():
    rng = np.random.default_rng(0)
    n = 1000
    g = [rng.standard_normal(n).astype(np.float32) for _ in range(3)]
    v0 = rng.standard_normal(n).astype(np.float32)
    a0 = rng.standard_normal(n).astype(np.float32)
    lr, mu = np.float32(0.01), np.float32(0.9)
    gm = ((g[0] + g[1]) + g[2]) / np.float32(3)
    v, _ = O.dense_apply(O.SGD, v0.copy(), None, g, 3.0, lr, mu)
    np.testing.assert_array_equal(v, v0 - gm * lr)
    v, a = O.dense_apply(O.MOMENTUM, v0.copy(), a0.copy(), g, 3.0, lr, mu)
    an = a0 * mu + gm
    np.testing.assert_array_equal(a, an)
    np.testing.assert_array_equal(v, v0 - an * lr)
    v, a = O.dense_apply(O.NESTEROV, v0.copy(), a0.copy(), g, 3.0, lr, mu)
    np.testing.assert_array_equal(v, v0 - (gm * lr + (an * mu) * lr))


def test_mean_of_identical_replicas_is_the_single_replica_step():
    rng = np.random.default_rng(1)
    g = rng.standard_normal(257).astype(np.float32)
    v0 = rng.standard_normal(257).astype(np.float32)
    for n in (1, 2):          # g + g and / 2 are exact; 3g already rounds in a sequential sum
        v, _ = O.dense_apply(O.SGD, v0.copy(), None, [g] * n, float(n), 0.01, 0.0)
        w, _ = O.dense_apply(O.SGD, v0.copy(), None, [g], 1.0, 0.01, 0.0)
        np.testing.assert_array_equal(v, w)


@settings(max_examples=200, deadline=None)
@given(vocab=st.integers(1, 5000), S=st.integers(1, 16), strat=st.sampled_from([O.MOD, O.DIV]))
def test_partition_is_a_bijection(vocab, S, strat):
    seen = set()
    rows = [O.shard_rows(vocab, S, p, strat) for p in range(S)]
    assert sum(rows) == vocab
    for id_ in range(0, vocab, max(1, vocab // 64)):
        sh, row = O.locate(vocab, S, strat, id_)
        assert 0 <= sh < S and 0 <= row < rows[sh]
        assert (sh, row) not in seen
        seen.add((sh, row))
    if strat == O.DIV:      # contiguous ranges, first vocab % S shards get one extra row
        assert rows == [vocab // S + (1 if p < vocab % S else 0) for p in range(S)]
        assert O.locate(vocab, S, strat, 0) == (0, 0)
    else:
        assert O.locate(vocab, S, strat, vocab - 1) == ((vocab - 1) % S, (vocab - 1) // S)


def test_gather_scatter_roundtrip_and_duplicates():
    rng = np.random.default_rng(2)
    vocab, dim, S = 103, 8, 4
    full = rng.standard_normal((vocab, dim)).astype(np.float32)
    ids = rng.integers(0, vocab, size=500).astype(np.int64)
    ids[:50] = 7                                        # heavy duplicates
    grads = rng.standard_normal((500, dim)).astype(np.float32)
    for strat in (O.MOD, O.DIV):
        shards = []
        for p in range(S):
            rows = [i for i in range(vocab) if O.locate(vocab, S, strat, i)[0] == p]
            rows.sort(key=lambda i: O.locate(vocab, S, strat, i)[1])
            shards.append(full[rows].copy())
        out = np.full((500, dim), np.nan, dtype=np.float32)
        for p in range(S):
            assert O.sparse_gather(shards[p], vocab, S, p, strat, ids, out) == 0
        np.testing.assert_array_equal(out, full[ids])   # every row filled by exactly one shard
        want = full.copy()
        for i, id_ in enumerate(ids):                   # the sequential definition
            want[id_] = want[id_] - np.float32(0.5) * grads[i]
        for p in range(S):
            O.sparse_scatter_sub(shards[p], vocab, S, p, strat, ids, grads, 0.5)
        for id_ in range(vocab):
            sh, row = O.locate(vocab, S, strat, id_)
            np.testing.assert_array_equal(shards[sh][row], want[id_])


def test_model_reference_worker_count_invariance():
    """an N-replica mean-of-means step equals a 1-replica step on the
    concatenated batch (fp64, so only rounding separates them)."""
    import torch
    from k8s_b200.syncps import mnist_init, synthetic_batch
    from oracle import model_ref as M
    p = [w.astype(np.float64) for w in mnist_init(0)]
    b = [synthetic_batch(r, 4) for r in range(2)]
    _, g0, _ = M.loss_and_grads(p, b[0][0], b[0][1], torch.float64)
    _, g1, _ = M.loss_and_grads(p, b[1][0], b[1][1], torch.float64)
    xc, yc = np.concatenate([b[0][0], b[1][0]]), np.concatenate([b[0][1], b[1][1]])
    _, gc, _ = M.loss_and_grads(p, xc, yc, torch.float64)
    for a, c, d in zip(g0, g1, gc):
        np.testing.assert_allclose((a + c) / 2, d, rtol=1e-9, atol=1e-12)
