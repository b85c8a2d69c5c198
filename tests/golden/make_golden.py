"""Generates tests/golden/*.json by importing the REFERENCE's Python replica programs from
/root/reference in this container (they cannot travel to the GPU box).

TensorFlow is absent, so the two modules are imported against a small stand-in `tensorflow`
package that implements only what their own code paths touch:
  * grpc_tensorflow_server.parse_cluster_spec needs a ClusterDef-like object with
    `job.add()` -> object with `.name` and a `.tasks` map  (pure string handling; the
    reference's code runs unmodified);
  * tf_smoke.run needs tf.Graph/device/constant/multiply/Session; the stand-in evaluates
    constant/multiply with numpy int32 and records the device string in force, so the
    golden output is "what the reference program asks each device to compute".
Run:  python tests/golden/make_golden.py     (writes next to this file)
"""
import contextlib
import importlib.util
import json
import os
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


# ---- stand-in tensorflow ---------------------------------------------------------------
class _Job:
    def __init__(self):
        self.name = ""
        self.tasks = {}


class _Jobs(list):
    def add(self):
        j = _Job()
        self.append(j)
        return j


class ClusterDef:
    def __init__(self):
        self.job = _Jobs()


class _Tensor:
    def __init__(self, value, device):
        self.value, self.device = value, device


class _TF(types.ModuleType):
    def __init__(self):
        super().__init__("tensorflow")
        self.__version__ = "stand-in"
        self.__git_version__ = "stand-in"
        self._device = [""]
        self.fetched = []

    class _G:
        @contextlib.contextmanager
        def as_default(self):
            yield self

    def Graph(self):
        return _TF._G()

    @contextlib.contextmanager
    def device(self, d):
        self._device.append(d if isinstance(d, str) else self._device[-1])
        try:
            yield
        finally:
            self._device.pop()

    def constant(self, value, shape=None):
        a = np.array(list(value), dtype=np.int32)
        return _Tensor(a.reshape(shape) if shape is not None else a, self._device[-1])

    def multiply(self, a, b):
        return _Tensor(a.value * b.value, self._device[-1])

    def global_variables_initializer(self):
        return _Tensor(None, self._device[-1])

    def ConfigProto(self, **kw):
        return kw

    def Session(self, target, config=None):
        tf = self

        class S:
            def __enter__(self):
                return self

            def __exit__(self, *a):
                return False

            def run(self, t):
                if t.value is not None:
                    tf.fetched.append((t.device, t.value))
                return t.value
        return S()


def install_stub():
    tf = _TF()
    sys.modules["tensorflow"] = tf
    for name in ["tensorflow.core", "tensorflow.core.protobuf", "tensorflow.python",
                 "tensorflow.python.platform", "tensorflow.python.training"]:
        sys.modules[name] = types.ModuleType(name)
    sys.modules["tensorflow.core.protobuf"].config_pb2 = types.ModuleType("config_pb2")
    sys.modules["tensorflow.core.protobuf"].tensorflow_server_pb2 = types.ModuleType("tsp")
    sys.modules["tensorflow.python.platform"].app = types.ModuleType("app")
    sys.modules["tensorflow.python.training"].server_lib = types.ModuleType("server_lib")
    return tf


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    tf = install_stub()
    gts = load(os.path.join(REF, "grpc_tensorflow_server/grpc_tensorflow_server.py"), "ref_gts")
    cases = [
        "master|master-0:2222,ps|localhost:2222;ps-1:2222,worker|worker-0:2222;worker-1:2222",
        "local|localhost:2222;localhost:2223",
        "ps|ps0:2222;ps1:2222",
        "master|example-job-master-76no-0:2222,ps|example-job-ps-76no-0:2222;example-job-ps-76no-1:2222,"
        "worker|example-job-worker-76no-0:2222",
        "", "nojobsep", "a|b|c", "|h:1", "ps|", "ps|a:1;;b:2", "ps|a:1,", "ps|a:1,,w|b:2", "w|x:1;",
    ]
    out = []
    for c in cases:
        cl = ClusterDef()
        try:
            gts.parse_cluster_spec(c, cl)
            out.append({"in": c, "jobs": [[j.name, {str(k): v for k, v in j.tasks.items()}] for j in cl.job]})
        except ValueError as e:
            out.append({"in": c, "error": str(e)})
    json.dump(out, open(os.path.join(HERE, "parse_cluster_spec.json"), "w"), indent=1)

    smoke = load(os.path.join(REF, "examples/tf_sample/tf_sample/tf_smoke.py"), "ref_smoke")
    clusters = {
        "tf_job.yaml": {"master": ["example-job-master-76no-0:2222"],
                        "worker": ["example-job-worker-76no-0:2222"],
                        "ps": ["example-job-ps-76no-0:2222", "example-job-ps-76no-1:2222"]},
        "tf_job_gpu.yaml": {"master": ["tf-smoke-gpu-master-abcd-0:2222"]},
        "unset": {},
    }
    sm = {}
    for name, cs in clusters.items():
        tf.fetched.clear()
        smoke.run(server=None, cluster_spec=cs)
        sm[name] = {"cluster": cs, "results": {d: v.tolist() for d, v in tf.fetched}}
    json.dump(sm, open(os.path.join(HERE, "tf_smoke.json"), "w"))
    print("wrote parse_cluster_spec.json (%d cases), tf_smoke.json (%d clusters)" % (len(out), len(sm)))


if __name__ == "__main__":
    main()
