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


@pytest.mark.parametrize("vocab,n_shards,shard,ids_kind", [
    ((1 << 23) + 5, 1, 0, "clustered"),      # 24-bit row keys: three radix passes of 8 bits
    ((1 << 22) + 3, 1, 0, "clustered"),      # 23 bits: three passes
    ((1 << 21) - 7, 1, 0, "clustered"),      # 21 bits: two passes of 11 bits (2048 bins, 64 KB of shared memory)
    (3 * (1 << 21), 3, 1, "clustered"),      # owned-id compaction in front of a two-pass sort
    (2000, 1, 0, "clustered"),               # one pass
])
def test_scatter_sub_key_widths(lib, vocab, n_shards, shard, ids_kind):
    """Every digit plan of the radix sort (sparse.cu: passes of at most 11 bits chosen from the shard's row count),
    with duplicates that differ only in their high key bits and out-of-range ids mixed in: bit-exact."""
    rng = np.random.default_rng(vocab)
    dim, n = 4, 30000
    base = rng.integers(0, vocab, size=40).astype(np.int64)
    ids = np.concatenate([base[rng.integers(0, 40, size=n // 2)],                       # heavy duplicates
                          (rng.integers(0, 8, size=n // 4) * (vocab // 8)).astype(np.int64),   # differ in the top bits only
                          rng.integers(0, vocab, size=n - n // 2 - n // 4).astype(np.int64)])
    ids[::997] = vocab + 3                                                              # out of range: skipped
    rng.shuffle(ids)
    grads = rng.standard_normal((n, dim)).astype(np.float32)
    rows = lib.sparse_shard_rows(vocab, n_shards, shard, L.PART_MOD)
    table = rng.standard_normal((rows, dim)).astype(np.float32)
    d_table, d_ids, d_grads = dev(table), dev(ids), dev(grads)
    scratch = torch.empty(lib.sparse_scratch_bytes(n), dtype=torch.uint8, device="cuda")
    lib.check(lib.sparse_scatter_sub(d_table.data_ptr(), vocab, dim, n_shards, shard, L.PART_MOD, d_ids.data_ptr(), n,
                                     d_grads.data_ptr(), 0.05, scratch.data_ptr(), scratch.numel(), 0))
    torch.cuda.synchronize()
    O.sparse_scatter_sub(table, vocab, n_shards, shard, L.PART_MOD, ids, grads, 0.05)
    np.testing.assert_array_equal(d_table.cpu().numpy().view(np.uint32), table.view(np.uint32))


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


@pytest.mark.parametrize("strategy", ["mod", "div"])
@pytest.mark.parametrize("n_gpus,n_shards,vocab,dim,batch,dist", [
    (2, 2, 1000, 64, 512, "uniform"), (4, 8, 5003, 128, 1000, "zipf"), (4, 3, 300, 33, 257, "uniform"),
    (2, 5, 64, 16, 2048, "one"), (8, 8, 100000, 128, 4096, "zipf"),
])
def test_sharded_embedding_across_emulated_ranks(lib, strategy, n_gpus, n_shards, vocab, dim, batch, dist):
    """The sparse path across GPU ranks (SURVEY.md par. 8e), all ranks emulated on one device (peer
    pointers are then ordinary device pointers): each rank looks its own batch up in ONE launch that
    reads every owner's shard, and each shard owner applies scatter_sub over the concatenation of all
    ranks' (ids, rows) in (rank, position) order.  Both bit-exact against the C oracle run on the
    concatenated batch, twice (scratch reuse)."""
    import json

    from k8s_b200.sparse import ShardedEmbedding
    from k8s_b200.syncps import SyncPSJob
    cluster = {"master": ["m:1"], "ps": ["p%d:1" % i for i in range(n_shards)]}
    if n_gpus > 1:
        cluster["worker"] = ["w%d:1" % i for i in range(n_gpus - 1)]
    jobs = [SyncPSJob(json.dumps({"cluster": cluster}), r, n_gpus, cuda_device=0) for r in range(n_gpus)]
    embs = [ShardedEmbedding(j, vocab, dim, batch, strategy) for j in jobs]
    for e in embs:
        assert e.owned == [k for k in range(n_shards) if jobs[0].lib.ps_gpu(jobs[0].h, k) == e.rank]
        e.attach_local([o for o in embs if o is not e])
    rng = np.random.default_rng(vocab * 7 + batch)
    strat = {"mod": L.PART_MOD, "div": L.PART_DIV}[strategy]
    tables = {}
    for e in embs:
        for k in e.owned:
            rows = lib.sparse_shard_rows(vocab, n_shards, k, strat)
            tables[k] = rng.standard_normal((max(rows, 1), dim)).astype(np.float32)
            e.set_shard(k, tables[k])
    for rep in range(2):
        ids = [make_ids(rng, vocab, batch, dist) for _ in range(n_gpus)]
        grads = [rng.standard_normal((batch, dim)).astype(np.float32) for _ in range(n_gpus)]
        for r, e in enumerate(embs):
            e.stage_ids(ids[r])
            got = e.lookup().cpu().numpy()
            want = np.full((batch, dim), np.nan, dtype=np.float32)
            for k in range(n_shards):
                O.sparse_gather(tables[k], vocab, n_shards, k, strat, ids[r], want)
            np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))
            assert int(e.bad.item()) == 0
        for r, e in enumerate(embs):
            e.stage_rows(grads[r])
        torch.cuda.synchronize()
        for e in embs:
            e.apply_owned(0.05)
        torch.cuda.synchronize()
        all_ids, all_grads = np.concatenate(ids), np.concatenate(grads)
        for k in range(n_shards):
            O.sparse_scatter_sub(tables[k], vocab, n_shards, k, strat, all_ids, all_grads, 0.05)
        for e in embs:
            for k in e.owned:
                np.testing.assert_array_equal(e.get_shard(k).view(np.uint32), tables[k].view(np.uint32))
    for j in jobs:
        j.close()


def test_sharded_embedding_chunked_long_runs_is_deterministic_and_accurate(lib):
    """chunk_long_runs=True on a heavy-tailed batch across 4 emulated ranks: rows hit <= 64 times stay
    bit-exact, the hot rows are deterministic and as accurate as the sequential loop (vs float64)."""
    import json

    from k8s_b200.sparse import ShardedEmbedding
    from k8s_b200.syncps import SyncPSJob
    n_gpus, vocab, dim, batch = 4, 2000, 64, 8192
    cluster = {"master": ["m:1"], "ps": ["p%d:1" % i for i in range(n_gpus)], "worker": ["w%d:1" % i for i in range(n_gpus - 1)]}
    rng = np.random.default_rng(3)
    ids = [make_ids(rng, vocab, batch, "zipf") for _ in range(n_gpus)]
    grads = [rng.standard_normal((batch, dim)).astype(np.float32) for _ in range(n_gpus)]
    table0 = rng.standard_normal((vocab, dim)).astype(np.float32)
    all_ids, all_grads = np.concatenate(ids), np.concatenate(grads)
    want = table0.copy()
    O.sparse_scatter_sub(want, vocab, 1, 0, O.MOD, all_ids, all_grads, 0.05)
    exact = table0.astype(np.float64)
    np.subtract.at(exact, all_ids, (np.float32(0.05) * all_grads).astype(np.float64))
    outs = []
    for rep in range(2):
        jobs = [SyncPSJob(json.dumps({"cluster": cluster}), r, n_gpus, cuda_device=0) for r in range(n_gpus)]
        embs = [ShardedEmbedding(j, vocab, dim, batch, "mod", chunk_long_runs=True) for j in jobs]
        for e in embs:
            e.attach_local([o for o in embs if o is not e])
            for k in e.owned:
                e.set_shard(k, table0[k::n_gpus])
        for r, e in enumerate(embs):
            e.stage_ids(ids[r])
            e.stage_rows(grads[r])
        torch.cuda.synchronize()
        for e in embs:
            e.apply_owned(0.05)
        torch.cuda.synchronize()
        full = np.empty_like(table0)
        for e in embs:
            for k in e.owned:
                full[k::n_gpus] = e.get_shard(k)
        outs.append(full)
        for j in jobs:
            j.close()
    np.testing.assert_array_equal(outs[0].view(np.uint32), outs[1].view(np.uint32))
    mult = np.bincount(all_ids, minlength=vocab)
    short, long_ = mult <= 64, mult > 64
    assert long_.any()
    np.testing.assert_array_equal(outs[0][short].view(np.uint32), want[short].view(np.uint32))
    err_gpu = np.abs(outs[0][long_] - exact[long_]).max(axis=1)
    err_seq = np.abs(want[long_] - exact[long_]).max(axis=1)
    scale = np.abs(exact[long_]).max(axis=1) + 1.0
    assert np.all(err_gpu <= 2.0 * err_seq + 4e-7 * scale), (err_gpu.max(), err_seq.max())
