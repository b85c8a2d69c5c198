"""Parity of the sparse embedding PS path (A15) with the C oracle: indices and values
bit-exact, duplicates included."""
import numpy as np
import pytest

from k8s_b200 import lib as L
from oracle import oracle as O

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def make_ids(rng, vocab, n, dist):
    if dist == "uniform":
        return rng.integers(0, vocab, size=n).astype(np.int64)
    if dist == "zipf":
        return np.minimum(rng.zipf(1.05, size=n) - 1, vocab - 1).astype(np.int64)
    if dist == "one":
        return np.full(n, vocab // 2, dtype=np.int64)
    raise ValueError(dist)


@pytest.mark.parametrize("strategy", [L.PART_MOD, L.PART_DIV])
@pytest.mark.parametrize("vocab,dim,n_shards,n,dist", [
    (1000, 64, 1, 4096, "uniform"), (1000, 64, 4, 4096, "zipf"), (1001, 128, 3, 5000, "uniform"),
    (50, 7, 16, 300, "uniform"), (100000, 128, 8, 65536, "zipf"), (300, 33, 2, 2049, "one"),
    (70000, 64, 1, 1, "uniform"), (5, 4, 8, 64, "uniform"),
])
def test_gather_and_scatter_sub_bit_exact(lib, strategy, vocab, dim, n_shards, n, dist):
    rng = np.random.default_rng(vocab + n)
    ids = make_ids(rng, vocab, n, dist)
    grads = rng.standard_normal((n, dim)).astype(np.float32)
    d_ids, d_grads = dev(ids), dev(grads)
    out_gpu = torch.full((n, dim), float("nan"), device="cuda")
    out_cpu = np.full((n, dim), np.nan, dtype=np.float32)
    scratch = torch.empty(lib.sparse_scratch_bytes(n), dtype=torch.uint8, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    for shard in range(n_shards):
        rows = lib.sparse_shard_rows(vocab, n_shards, shard, strategy)
        table = rng.standard_normal((max(rows, 1), dim)).astype(np.float32)
        d_table = dev(table)
        lib.check(lib.sparse_gather(d_table.data_ptr(), vocab, dim, n_shards, shard, strategy, d_ids.data_ptr(),
                                    n, out_gpu.data_ptr(), bad.data_ptr(), 0))
        O.sparse_gather(table, vocab, n_shards, shard, strategy, ids, out_cpu)
        for rep in range(2):     # twice: the scratch is reused and must not carry state over
            lib.check(lib.sparse_scatter_sub(d_table.data_ptr(), vocab, dim, n_shards, shard, strategy,
                                             d_ids.data_ptr(), n, d_grads.data_ptr(), 0.05,
                                             scratch.data_ptr(), scratch.numel(), 0))
            O.sparse_scatter_sub(table, vocab, n_shards, shard, strategy, ids, grads, 0.05)
        torch.cuda.synchronize()
        np.testing.assert_array_equal(d_table.cpu().numpy().view(np.uint32), table.view(np.uint32))
    np.testing.assert_array_equal(out_gpu.cpu().numpy().view(np.uint32), out_cpu.view(np.uint32))
    assert not np.isnan(out_cpu).any() and int(bad.item()) == 0


@pytest.mark.parametrize("vocab,dim,n,dist", [(1000, 64, 20000, "zipf"), (300, 33, 5000, "one"), (5000, 128, 65536, "zipf"),
                                              (100000, 128, 200000, "uniform")])
def test_scatter_sub_chunk_parallel_long_runs(lib, vocab, dim, n, dist):
    """With tfg_sparse_scratch_bytes_fast() of scratch, rows hit by more than 64 ids are summed
    chunk-parallel.  Rows hit <= 64 times stay bit-exact with the sequential oracle; the others
    must be deterministic and as accurate as the oracle is (both measured against float64)."""
    rng = np.random.default_rng(vocab + n)
    ids = make_ids(rng, vocab, n, dist)
    grads = rng.standard_normal((n, dim)).astype(np.float32)
    table0 = rng.standard_normal((vocab, dim)).astype(np.float32)
    want = table0.copy()
    O.sparse_scatter_sub(want, vocab, 1, 0, O.MOD, ids, grads, 0.05)
    exact = table0.astype(np.float64)
    np.subtract.at(exact, ids, (np.float32(0.05) * grads).astype(np.float64))
    d_ids, d_grads = dev(ids), dev(grads)
    scratch = torch.empty(lib.sparse_scratch_bytes_fast(n, dim), dtype=torch.uint8, device="cuda")
    assert scratch.numel() > lib.sparse_scratch_bytes(n)
    outs = []
    for rep in range(2):
        d_table = dev(table0)
        lib.check(lib.sparse_scatter_sub(d_table.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, d_ids.data_ptr(), n,
                                         d_grads.data_ptr(), 0.05, scratch.data_ptr(), scratch.numel(), 0))
        torch.cuda.synchronize()
        outs.append(d_table.cpu().numpy())
    np.testing.assert_array_equal(outs[0].view(np.uint32), outs[1].view(np.uint32))      # deterministic
    mult = np.bincount(ids, minlength=vocab)
    short, long_ = mult <= 64, mult > 64
    np.testing.assert_array_equal(outs[0][short].view(np.uint32), want[short].view(np.uint32))
    if dist != "uniform":
        assert long_.any()
    if long_.any():
        err_gpu = np.abs(outs[0][long_] - exact[long_]).max(axis=1)
        err_seq = np.abs(want[long_] - exact[long_]).max(axis=1)
        scale = np.abs(exact[long_]).max(axis=1) + 1.0
        assert np.all(err_gpu <= 2.0 * err_seq + 4e-7 * scale), (err_gpu.max(), err_seq.max())


def test_out_of_range_ids_are_counted_and_skipped(lib):
    vocab, dim = 100, 16
    ids = np.array([5, -1, 100, 99, 1 << 40, 5], dtype=np.int64)
    table = np.arange(vocab * dim, dtype=np.float32).reshape(vocab, dim)
    d_table, d_ids = dev(table), dev(ids)
    out = torch.zeros((len(ids), dim), device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    lib.check(lib.sparse_gather(d_table.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, d_ids.data_ptr(), len(ids),
                                out.data_ptr(), bad.data_ptr(), 0))
    assert int(bad.item()) == 3
    ref = np.zeros((len(ids), dim), dtype=np.float32)
    assert O.sparse_gather(table, vocab, 1, 0, O.MOD, ids, ref) == 3
    np.testing.assert_array_equal(out.cpu().numpy(), ref)
    g = np.ones((len(ids), dim), dtype=np.float32)
    scratch = torch.empty(lib.sparse_scratch_bytes(len(ids)), dtype=torch.uint8, device="cuda")
    lib.check(lib.sparse_scatter_sub(d_table.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, d_ids.data_ptr(), len(ids),
                                     dev(g).data_ptr(), 1.0, scratch.data_ptr(), scratch.numel(), 0))
    O.sparse_scatter_sub(table, vocab, 1, 0, O.MOD, ids, g, 1.0)
    np.testing.assert_array_equal(d_table.cpu().numpy(), table)
    # argument checking
    assert lib.sparse_scatter_sub(d_table.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, d_ids.data_ptr(), len(ids),
                                  dev(g).data_ptr(), 1.0, scratch.data_ptr(), 16, 0) == 2   # scratch too small
    assert lib.sparse_gather(d_table.data_ptr(), vocab, dim, 2, 2, L.PART_MOD, d_ids.data_ptr(), 1,
                             out.data_ptr(), None, 0) == 2
    lib.check(lib.sparse_gather(d_table.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, d_ids.data_ptr(), 0,
                                out.data_ptr(), None, 0))                                    # empty batch


def test_full_size_roundtrip_property(lib):
    """BASELINE config 5 size: 1M x 128 table, 1M ids.  gather . scatter_sub with unique ids
    and scale 1 zeroes exactly the gathered rows; a permutation of ids gives the same table."""
    vocab, dim, n = 1 << 20, 128, 1 << 20
    table = torch.randn(vocab, dim, device="cuda")
    keep = table.clone()
    perm = torch.randperm(vocab, device="cuda")[:n].contiguous()
    rows = torch.empty(n, dim, device="cuda")
    lib.check(lib.sparse_gather(table.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, perm.data_ptr(), n, rows.data_ptr(), None, 0))
    assert torch.equal(rows, keep[perm])
    scratch = torch.empty(lib.sparse_scratch_bytes(n), dtype=torch.uint8, device="cuda")
    lib.check(lib.sparse_scatter_sub(table.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, perm.data_ptr(), n,
                                     rows.data_ptr(), 1.0, scratch.data_ptr(), scratch.numel(), 0))
    torch.cuda.synchronize()
    assert not table[perm].any()
    assert torch.count_nonzero(table.abs().sum(1)) == vocab - n
