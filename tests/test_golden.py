"""Golden vectors produced by the reference's own Python replica programs (imported from
/root/reference by tests/golden/make_golden.py against a stand-in tensorflow) versus
(a) the host-side mirrors, (b) the C ABI, (c) the C oracle."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from k8s_b200 import grpc_tensorflow_server as G
from k8s_b200 import tf_smoke
from oracle import oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")
PCS = json.load(open(os.path.join(GOLD, "parse_cluster_spec.json")))
SMOKE = json.load(open(os.path.join(GOLD, "tf_smoke.json")))


@pytest.mark.parametrize("case", PCS, ids=[repr(c["in"])[:40] for c in PCS])
def test_parse_cluster_spec_python_mirror(case):
    if "error" in case:
        with pytest.raises(ValueError) as e:
            G.parse_cluster_spec(case["in"])
        assert str(e.value) == case["error"]
    else:
        got = G.parse_cluster_spec(case["in"])
        assert [[k, {str(i): a for i, a in v.items()}] for k, v in got.items()] == case["jobs"]


@pytest.mark.parametrize("case", PCS, ids=[repr(c["in"])[:40] for c in PCS])
def test_parse_cluster_spec_c_abi(case, lib):
    if "error" in case:
        with pytest.raises(ValueError) as e:
            lib.parse_cluster_spec_json(case["in"])
        assert str(e.value) == case["error"]
    else:
        got = json.loads(lib.parse_cluster_spec_json(case["in"]))
        want = {name: [tasks[str(i)] for i in range(len(tasks))] for name, tasks in case["jobs"]}
        assert got == want and list(got) == [n for n, _ in case["jobs"]]


def test_parse_is_inverse_of_transform():
    from k8s_b200.trainer import transform_cluster_spec_for_default_ps as tx
    cs = {"master": ["m-0:1"], "ps": ["p-0:2", "p-1:2"], "worker": ["w-0:3"]}
    assert G.cluster_lists(G.parse_cluster_spec(tx(cs))) == cs


@pytest.mark.parametrize("name", sorted(SMOKE))
def test_smoke_oracle_matches_reference_program(name):
    g = SMOKE[name]
    n_tasks = sum(len(v) for v in g["cluster"].values())
    assert len(g["results"]) == n_tasks
    want = O.smoke(max(n_tasks, 1))
    for dev, val in g["results"].items():
        np.testing.assert_array_equal(np.array(val, dtype=np.int32), want[0])
    # sum 328350, max 9801
    assert int(want[0].sum()) == 328350 and int(want[0].max()) == 9801
    assert want[0][3][7] == 37 * 37


def test_tf_smoke_mirror_role_dispatch(monkeypatch):
    cfg = {"cluster": SMOKE["tf_job.yaml"]["cluster"], "environment": "cloud"}
    for typ in ("ps", "worker"):
        monkeypatch.setenv("TF_CONFIG", json.dumps(dict(cfg, task={"type": typ, "index": 0})))
        assert tf_smoke.main([]) is None          # serve-only roles, tf_smoke.py:126-132
    monkeypatch.setenv("TF_CONFIG", json.dumps(dict(cfg, task={"type": "evaluator", "index": 0})))
    with pytest.raises(ValueError, match="invalid job_type evaluator"):
        tf_smoke.main([])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(SMOKE))
def test_smoke_cuda_matches_reference_program(name, monkeypatch):
    g = SMOKE[name]
    monkeypatch.setenv("TF_CONFIG", json.dumps({"cluster": g["cluster"], "task": {"type": "master", "index": 0},
                                                "environment": "cloud"}))
    got = tf_smoke.main([])
    assert set(got) == set(g["results"])
    for dev, val in g["results"].items():
        np.testing.assert_array_equal(got[dev], np.array(val, dtype=np.int32))
