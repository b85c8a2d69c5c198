"""The C-ABI library loads and exports every symbol include/tfgraft.h declares; host-only
entry points (cluster, placement) are exercised without a GPU, compute ones are not."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from k8s_b200 import lib as L
from k8s_b200.cluster import RankMap, round_robin_placement
from k8s_b200.syncps import MNIST_SHAPES, MNIST_VARS, SyncPSJob
from oracle import oracle as O

TF_JOB_YAML = {"cluster": {"master": ["example-job-master-76no-0:2222"],
                           "ps": ["example-job-ps-76no-0:2222", "example-job-ps-76no-1:2222"],
                           "worker": ["example-job-worker-76no-0:2222"]},
               "task": {"type": "master", "index": 0}, "environment": "cloud"}


def test_library_exports_every_declared_symbol(lib):
    declared = L.declared_symbols()
    assert len(declared) >= 40
    for name in declared:
        assert hasattr(lib.dll, name), "libtfgraft.so does not export %s" % name
    assert sorted(L._SIGS) == declared, "lib.py signatures and include/tfgraft.h disagree"
    assert lib.version() == 100


def test_job_create_parses_tf_config_and_maps_ranks():
    for n_gpus in (1, 2, 8):
        rm = RankMap(TF_JOB_YAML["cluster"], n_gpus)
        for rank in range(n_gpus):
            j = SyncPSJob(json.dumps(TF_JOB_YAML), rank, n_gpus)
            assert (j.n_replicas, j.n_ps) == (2, 2) == (rm.n_replicas, rm.n_ps)
            assert (j.num_tasks("master"), j.num_tasks("worker"), j.num_tasks("ps"), j.num_tasks("x")) == (1, 1, 2, -1)
            assert j.local_replicas == rm.local_replicas(rank)
            for r, (job, t) in enumerate(rm.replicas()):
                assert j.lib.replica_gpu(j.h, r) == rm.gpu_of(job, t)
            for k in range(2):
                assert j.lib.ps_gpu(j.h, k) == rm.gpu_of("ps", k)
            j.close()


def test_job_create_rejects_bad_input(lib):
    h = C.c_void_p()
    assert lib.job_create(b"{not json", 0, 1, C.byref(h)) == 3          # permanent (1..127)
    assert "TF_CONFIG" in lib.error(None)
    assert lib.job_create(b'{"cluster": {"ps": "x"}}', 0, 1, C.byref(h)) == 3
    assert lib.job_create(b"{}", 2, 2, C.byref(h)) == 2
    assert lib.job_create(b"{}", 0, 99, C.byref(h)) == 2
    j = SyncPSJob("{}", 0, 1)                                           # TF_CONFIG unset: local case
    assert (j.n_replicas, j.n_ps, j.local_replicas) == (1, 0, [0])


def test_round_robin_placement_matches_oracle_and_survey_numbers():
    """MNIST CNN over 2 PS -> loads [3273504, 1130]; over 8 PS ->
    [800, 32, 51200, 64, 3211264, 1024, 10240, 10]."""
    for n_ps, want in [(2, [3273504, 1130]), (8, [800, 32, 51200, 64, 3211264, 1024, 10240, 10]), (3, None), (1, None)]:
        cluster = {"master": ["m:1"], "ps": ["p%d:1" % i for i in range(n_ps)]}
        j = SyncPSJob(json.dumps({"cluster": cluster}), 0, 1)
        tasks = []
        for name, shape in zip(MNIST_VARS, MNIST_SHAPES):
            _, t = j.register(name, shape)
            tasks.append(t)
        _, t = j.register("global_step", (), dtype=L.DT_I64, trainable=False)
        tasks.append(t)
        _, t = j.register("some_op_result", (4,), op_type="Identity")
        assert t == -1                                                   # not a PS op: stays on the worker
        ops = ["VariableV2"] * 9
        assert tasks == O.round_robin(ops, n_ps) == round_robin_placement(ops, n_ps)
        loads = [0] * n_ps
        for (name, shape), t in zip(zip(MNIST_VARS, MNIST_SHAPES), tasks):
            loads[t] += int(np.prod(shape))
        if want:
            assert loads == want
        assert sum(loads) == 3274634
    j = SyncPSJob(json.dumps({"cluster": {"master": ["m:1"]}}), 0, 1)    # no ps job: all local
    assert j.register("w", (3, 3))[1] == -1


def test_sparse_shard_rows_matches_oracle(lib):
    for vocab in (1, 7, 10, 1000, 1001):
        for S in (1, 2, 3, 8, 16):
            for strat in (L.PART_MOD, L.PART_DIV):
                rows = [lib.sparse_shard_rows(vocab, S, p, strat) for p in range(S)]
                assert rows == [O.shard_rows(vocab, S, p, strat) for p in range(S)]
                assert sum(rows) == vocab
    assert lib.sparse_shard_rows(10, 0, 0, 0) == -1
    assert lib.sparse_scratch_bytes(0) >= 256 and lib.sparse_scratch_bytes(1 << 20) >= 16 << 20
    # the array behind tfg_sparse_scatter_sub_static is exactly what the sort asks for at its id limit, and the
    # Python mirror of the header constants agrees with the header
    assert lib.sparse_scratch_bytes(L.SPARSE_STATIC_MAX_IDS) == L.SPARSE_STATIC_SCRATCH_BYTES
    header = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "tfgraft.h")).read()
    assert "#define TFG_SPARSE_STATIC_MAX_IDS %d" % L.SPARSE_STATIC_MAX_IDS in header
    assert "#define TFG_SPARSE_STATIC_SCRATCH_BYTES %d" % L.SPARSE_STATIC_SCRATCH_BYTES in header


def test_compute_entry_points_fail_loudly_without_gpu(lib):
    if lib.device_count_or_zero() > 0:
        pytest.skip("a GPU is present")
    j = SyncPSJob(json.dumps(TF_JOB_YAML), 0, 1)
    with pytest.raises(L.TfgError) as e:
        j.build_mnist(8)
    assert e.value.retryable and "CUDA" in e.value.message          # >= 128: no silent CPU path
