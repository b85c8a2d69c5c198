"""tfg_sparse_scatter_sub_static -- the scatter-sub variant for callers that cannot allocate
device memory (a PS process whose GPU memory a framework allocator already owns, or a step
loop that must avoid cudaMalloc's device-wide synchronisation).  Its sort scratch is a
fixed array inside the library image, so this test hands it nothing but the table, the ids
and the gradient rows, and checks the result against tfg_sparse_scatter_sub fed a scratch
buffer we allocate ourselves.  Both paths apply duplicate ids in order of appearance, so
the comparison is bit-for-bit (the tolerance every sparse test here uses: none).

Two calls are made back to back on purpose: the built-in scratch is shared by all calls and
still holds the first call's sort when the second starts, which must not matter.

The file name sorts last so that a run stopping at the first failure has already run every
other test; it also has to pass on its own in a fresh process (first use of the scratch).
"""
import numpy as np
import pytest

from k8s_b200 import lib as L

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_static_scratch_matches_caller_scratch_on_two_calls_in_a_row(lib):
    vocab, dim, n_shards, shard = 5000, 64, 2, 1
    rng = np.random.default_rng(2024)
    rows_in_shard = lib.sparse_shard_rows(vocab, n_shards, shard, L.PART_MOD)
    table0 = rng.standard_normal((rows_in_shard, dim)).astype(np.float32)
    want, got = dev(table0), dev(table0)
    # call 1: a full-size batch with heavy duplicates; call 2: a smaller, different batch
    for call, n in enumerate([L.SPARSE_STATIC_MAX_IDS, 777]):
        ids = rng.integers(0, vocab, size=n).astype(np.int64)
        ids[: n // 8] = 4321                                  # one hot row, applied in order
        grads = rng.standard_normal((n, dim)).astype(np.float32)
        d_ids, d_grads = dev(ids), dev(grads)
        scratch = torch.empty(lib.sparse_scratch_bytes(n), dtype=torch.uint8, device="cuda")
        lib.check(lib.sparse_scatter_sub(want.data_ptr(), vocab, dim, n_shards, shard, L.PART_MOD,
                                         d_ids.data_ptr(), n, d_grads.data_ptr(), 0.05,
                                         scratch.data_ptr(), scratch.numel(), 0))
        lib.check(lib.sparse_scatter_sub_static(got.data_ptr(), vocab, dim, n_shards, shard, L.PART_MOD,
                                                d_ids.data_ptr(), n, d_grads.data_ptr(), 0.05, 0))
        torch.cuda.synchronize()
        a, b = want.cpu().numpy().view(np.uint32), got.cpu().numpy().view(np.uint32)
        np.testing.assert_array_equal(b, a, err_msg="call %d differs from tfg_sparse_scatter_sub" % (call + 1))
        assert not np.array_equal(want.cpu().numpy(), table0)  # the update really happened


def test_static_scratch_call_allocates_no_device_memory_and_takes_odd_row_widths(lib):
    """What the entry point is for: once the first call on the device has looked up the built-in
    array, a call must leave the amount of free device memory exactly where it was (no cudaMalloc
    behind the caller's back).  Row width 33 is not a multiple of 4: the scalar load/store path."""
    vocab, dim, n = 700, 33, 5000
    rng = np.random.default_rng(7)
    table0 = rng.standard_normal((vocab, dim)).astype(np.float32)
    ids = rng.integers(0, vocab, size=n).astype(np.int64)
    grads = rng.standard_normal((n, dim)).astype(np.float32)
    want, got, d_ids, d_grads = dev(table0), dev(table0), dev(ids), dev(grads)
    scratch = torch.empty(lib.sparse_scratch_bytes(n), dtype=torch.uint8, device="cuda")
    lib.check(lib.sparse_scatter_sub(want.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, d_ids.data_ptr(), n,
                                     d_grads.data_ptr(), 0.05, scratch.data_ptr(), scratch.numel(), 0))
    lib.check(lib.sparse_scatter_sub_static(got.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, d_ids.data_ptr(), n,
                                            d_grads.data_ptr(), 0.05, 0))      # first use on this device may set up
    torch.cuda.synchronize()
    np.testing.assert_array_equal(got.cpu().numpy().view(np.uint32), want.cpu().numpy().view(np.uint32))
    free_before = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        lib.check(lib.sparse_scatter_sub_static(got.data_ptr(), vocab, dim, 1, 0, L.PART_MOD, d_ids.data_ptr(), n,
                                                d_grads.data_ptr(), 0.05, 0))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before


def test_static_scratch_rejects_oversized_batches_and_empty_is_a_noop(lib):
    table = torch.zeros((10, 4), device="cuda")
    ids = torch.zeros(1, dtype=torch.int64, device="cuda")
    rows = torch.ones((1, 4), device="cuda")
    rc = lib.sparse_scatter_sub_static(table.data_ptr(), 10, 4, 1, 0, L.PART_MOD, ids.data_ptr(),
                                       L.SPARSE_STATIC_MAX_IDS + 1, rows.data_ptr(), 1.0, 0)
    assert rc == 5                                            # TFG_ELIMIT, a permanent error
    msg = lib.error(None)
    assert "TFG_SPARSE_STATIC_MAX_IDS" in msg and str(L.SPARSE_STATIC_MAX_IDS + 1) in msg
    lib.check(lib.sparse_scatter_sub_static(table.data_ptr(), 10, 4, 1, 0, L.PART_MOD, ids.data_ptr(), 0,
                                            rows.data_ptr(), 1.0, 0))
    torch.cuda.synchronize()
    assert not table.any()
