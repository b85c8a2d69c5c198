"""Parity of the dense PS update kernels (A12-A14) with the C oracle: bit-exact."""
import ctypes as C
import json

import numpy as np
import pytest

from k8s_b200 import lib as L
from k8s_b200.syncps import SyncPSJob
from oracle import oracle as O

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("kind", ["sgd", "momentum", "nesterov"])
@pytest.mark.parametrize("count", [1, 3, 4, 1000, 4096, 4097, 100003, (1 << 20) + 5])
@pytest.mark.parametrize("n_grads", [1, 3, 8])
def test_dense_apply_bit_exact(lib, kind, count, n_grads):
    rng = np.random.default_rng(count * 31 + n_grads)
    v0 = rng.standard_normal(count).astype(np.float32)
    a0 = rng.standard_normal(count).astype(np.float32)
    gs = [rng.standard_normal(count).astype(np.float32) for _ in range(n_grads)]
    gs[0][:1] = np.float32(1e-38)         # a denormal-range product on the way
    lr, mu = 0.01, 0.9
    want_v, want_a = O.dense_apply(O.KINDS[kind], v0.copy(), a0.copy(), gs, float(n_grads), lr, mu)
    dv, da, dg = dev(v0), dev(a0), [dev(g) for g in gs]
    pulls = [torch.empty_like(dv) for _ in range(2)]
    rc = lib.dense_apply(L.OPT_KINDS[kind], dv.data_ptr(), da.data_ptr(), L.ptr_array(dg), n_grads,
                         float(n_grads), count, lr, mu, L.ptr_array(pulls), 2, 0)
    lib.check(rc)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(bits(dv.cpu().numpy()), bits(want_v))
    for p in pulls:
        np.testing.assert_array_equal(bits(p.cpu().numpy()), bits(want_v))
    if kind != "sgd":
        np.testing.assert_array_equal(bits(da.cpu().numpy()), bits(want_a))
    else:
        np.testing.assert_array_equal(da.cpu().numpy(), a0)          # SGD leaves the slot alone


def test_dense_apply_unaligned_and_errors(lib):
    rng = np.random.default_rng(5)
    n = 1001
    base_v, base_g = dev(rng.standard_normal(n + 1).astype(np.float32)), dev(rng.standard_normal(n + 1).astype(np.float32))
    v0, g0 = base_v.cpu().numpy()[1:].copy(), base_g.cpu().numpy()[1:].copy()
    want, _ = O.dense_apply(O.SGD, v0.copy(), None, [g0], 1.0, 0.1, 0.0)
    rc = lib.dense_apply(L.OPT_SGD, base_v.data_ptr() + 4, None, L.ptr_array([base_g.data_ptr() + 4]), 1, 1.0,
                         n, 0.1, 0.0, None, 0, 0)
    lib.check(rc)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(bits(base_v.cpu().numpy()[1:]), bits(want))
    assert lib.dense_apply(L.OPT_MOMENTUM, base_v.data_ptr(), None, L.ptr_array([base_g]), 1, 1.0, n, 0.1, 0.9,
                           None, 0, 0) == 2
    assert lib.dense_apply(7, base_v.data_ptr(), None, L.ptr_array([base_g]), 1, 1.0, n, 0.1, 0.9, None, 0, 0) == 2
    assert lib.dense_apply(L.OPT_SGD, base_v.data_ptr(), None, L.ptr_array([base_g]), 65, 1.0, n, 0.1, 0.9,
                           None, 0, 0) == 5
    lib.check(lib.dense_apply(L.OPT_SGD, base_v.data_ptr(), None, L.ptr_array([base_g]), 1, 1.0, 0, 0.1, 0.9,
                              None, 0, 0))                            # empty shard: a no-op


def _cluster(n_rep, n_ps):
    c = {"master": ["m-0:2222"], "ps": ["p-%d:2222" % i for i in range(n_ps)]}
    if n_rep > 1:
        c["worker"] = ["w-%d:2222" % i for i in range(n_rep - 1)]
    return json.dumps({"cluster": c, "task": {"type": "master", "index": 0}, "environment": "cloud"})


@pytest.mark.parametrize("layout", ["logical", "balanced"])
@pytest.mark.parametrize("kind", ["sgd", "momentum", "nesterov"])
@pytest.mark.parametrize("n_gpus,n_rep,n_ps", [(1, 1, 1), (1, 2, 2), (2, 2, 2), (4, 4, 3), (4, 6, 8)])
def test_job_push_apply_pull_emulated_ranks(layout, kind, n_gpus, n_rep, n_ps):
    """All GPU ranks of a job emulated on one device as separate jobs attached to each
    other (peer pointers are then ordinary device pointers).  Every rank's parameter
    buffer must end up bit-identical to the oracle's, for both physical layouts."""
    shapes = [(37,), (5, 13), (1000,), (64, 65), (3,), (4096,), (2, 2, 2, 2)]
    jobs = [SyncPSJob(_cluster(n_rep, n_ps), r, n_gpus, cuda_device=0) for r in range(n_gpus)]
    rng = np.random.default_rng(n_gpus * 100 + n_rep * 10 + n_ps)
    init = [rng.standard_normal(s).astype(np.float32) for s in shapes]
    grads = [[rng.standard_normal(s).astype(np.float32) for s in shapes] for _ in range(n_rep)]
    for j in jobs:
        for i, s in enumerate(shapes):
            j.register("v%d" % i, s)
        j.finalize(layout)
        j.optimizer(kind, 0.05, 0.9)
        for i, w in enumerate(init):
            j.set_var(i, w)
        for slot, r in enumerate(j.local_replicas):
            for i in range(len(shapes)):
                j.set_grad(slot, i, grads[r][i])
    for a in jobs:
        for b in jobs:
            if a is not b:
                a.attach_local_peer(b.gpu_rank, b)
    want = [w.copy().reshape(-1) for w in init]
    acc = [np.zeros_like(w) for w in want]
    for step in range(2):
        for j in jobs:
            j.push_apply_pull(with_barriers=False)
        for j in jobs:
            j.sync()
        for i in range(len(shapes)):
            O.dense_apply(O.KINDS[kind], want[i], acc[i], [grads[r][i].reshape(-1) for r in range(n_rep)],
                          float(n_rep), 0.05, 0.9)
        for j in jobs:
            assert j.global_step == step + 1
            for i in range(len(shapes)):
                np.testing.assert_array_equal(bits(j.get_var(i)), bits(want[i]), err_msg="rank %d var %d" % (j.gpu_rank, i))
    # placement: the owner of each variable is its PS task's GPU (A11), on every rank alike
    tasks = O.round_robin(["VariableV2"] * len(shapes), n_ps)
    for j in jobs:
        for i in range(len(shapes)):
            p = j.placement(i)
            assert p["ps_task"] == tasks[i] and p["gpu"] == tasks[i] % n_gpus and p["count"] == int(np.prod(shapes[i]))
            assert p["offset"] % 32 == 0
    if kind != "sgd" and layout == "logical":
        for j in jobs:                       # the slot lives with the owner only
            for i in range(len(shapes)):
                got = j.get_accum(i)
                if j.placement(i)["gpu"] == j.gpu_rank:
                    np.testing.assert_array_equal(bits(got), bits(acc[i]))
                else:
                    assert not got.any()
    for j in jobs:
        j.close()


def test_apply_owned_from_presummed_gradient():
    """NCCL-baseline split: gradient already reduced (summed) into slot 0."""
    j = SyncPSJob(_cluster(4, 1), 0, 1, cuda_device=0)
    j.register("w", (5000,))
    j.finalize("balanced")
    j.optimizer("momentum", 0.1, 0.5)
    rng = np.random.default_rng(9)
    w, g = rng.standard_normal(5000).astype(np.float32), rng.standard_normal(5000).astype(np.float32)
    j.set_var(0, w)
    j.set_grad(0, 0, g)
    j.apply_owned(4)
    j.sync()
    want, acc = O.dense_apply(O.MOMENTUM, w.copy(), np.zeros_like(w), [g], 4.0, 0.1, 0.5)
    np.testing.assert_array_equal(bits(j.get_var(0)), bits(want))
    np.testing.assert_array_equal(bits(j.get_accum(0)), bits(acc))


def test_full_size_shard_properties(lib):
    """BASELINE config 5 size (64M parameters here; 1B in bench --sweep): size-independent
    properties instead of an element-wise oracle -- linearity in lr for SGD and the
    zero-gradient fixed point."""
    n = 64 << 20
    g = torch.randn(n, device="cuda")
    v = torch.randn(n, device="cuda")
    v0 = v.clone()
    z = torch.zeros_like(g)
    lib.check(lib.dense_apply(L.OPT_SGD, v.data_ptr(), None, L.ptr_array([z]), 1, 1.0, n, 0.5, 0.0, None, 0, 0))
    assert torch.equal(v, v0)                                           # g = 0 is a fixed point
    lib.check(lib.dense_apply(L.OPT_SGD, v.data_ptr(), None, L.ptr_array([g]), 1, 1.0, n, 0.25, 0.0, None, 0, 0))
    assert torch.equal(v, v0 - g * 0.25)                                # power-of-two lr: exact in torch too
    pull = torch.empty_like(v)
    lib.check(lib.dense_apply(L.OPT_SGD, v.data_ptr(), None, L.ptr_array([g, g]), 2, 2.0, n, 0.25, 0.0,
                              L.ptr_array([pull]), 1, 0))
    assert torch.equal(pull, v) and torch.equal(v, (v0 - g * 0.25) - g * 0.25)
