"""Host-side handle of one GPU rank's share of a synchronous PS job.

This is what a MASTER/WORKER replica process (k8s_b200.tf_smoke, bench.py) holds in place
of ``tf.train.Server`` + ``tf.Session`` (examples/tf_sample/tf_sample/tf_smoke.py:101-112,
71-76): it is built from the replica's TF_CONFIG, owns the parameter shards of the PS
tasks mapped to this GPU, and runs the hosted training replicas.  All arithmetic happens
in libtfgraft; torch is used for pinned host memory, streams/events and the
torch.distributed exchange of peer handles.
"""
from __future__ import annotations

import ctypes as C
import json
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import lib as L

MNIST_VARS = ["W_conv1", "b_conv1", "W_conv2", "b_conv2", "W_fc1", "b_fc1", "W_fc2", "b_fc2"]
MNIST_SHAPES = [(5, 5, 1, 32), (32,), (5, 5, 32, 64), (64,), (3136, 1024), (1024,), (1024, 10), (10,)]


def mnist_init(seed: int = 0) -> List[np.ndarray]:
    """Weights of the sample model: truncated normal (|z| <= 2), sigma 0.1; biases 0.1
    (SURVEY.md §8d).  Generated on the CPU from a fixed seed and uploaded, never by a
    GPU-side RNG, so every rank and the oracle start from identical bits."""
    rng = np.random.default_rng(seed)
    out = []
    for shape in MNIST_SHAPES:
        if len(shape) == 1:
            out.append(np.full(shape, 0.1, dtype=np.float32))
            continue
        z = rng.standard_normal(size=shape)
        bad = np.abs(z) > 2.0
        while bad.any():
            z[bad] = rng.standard_normal(size=int(bad.sum()))
            bad = np.abs(z) > 2.0
        out.append((0.1 * z).astype(np.float32))
    return out


def synthetic_batch(replica: int, batch: int, step: int = 0):
    """Images N(0,1) fp32 [B,28,28,1] and labels uniform in [0,10), seed 1000+replica
    (SURVEY.md §8d); ``step`` offsets the stream so successive steps differ."""
    rng = np.random.default_rng([1000 + replica, step])
    x = rng.standard_normal(size=(batch, 28, 28, 1)).astype(np.float32)
    y = rng.integers(0, 10, size=(batch,)).astype(np.int32)
    return x, y


def resnet50_init(job: "SyncPSJob", seed: int = 0) -> Dict[int, np.ndarray]:
    """{variable index: initial value} for every trainable variable of a built ResNet-50."""
    rng = np.random.default_rng(seed)
    out: Dict[int, np.ndarray] = {}
    for v in range(job.n_vars):
        name, shape, trainable = job.var_info(v)
        if not trainable:
            continue
        if name.endswith("/weights") and len(shape) == 4:
            fan_in = shape[0] * shape[1] * shape[2]
            out[v] = (rng.standard_normal(size=shape) * np.sqrt(2.0 / fan_in)).astype(np.float32)
        elif name == "fc/weights":
            out[v] = (rng.standard_normal(size=shape) * 0.01).astype(np.float32)
        elif name.endswith("/gamma"):
            out[v] = np.ones(shape, dtype=np.float32)
        else:
            out[v] = np.zeros(shape, dtype=np.float32)
    return out


def synthetic_images(replica: int, batch: int, hw: int, classes: int, step: int = 0):
    """Synthetic ImageNet-shaped batch: N(0,1) fp32 [B,hw,hw,3], labels uniform in [0, classes)."""
    rng = np.random.default_rng([2000 + replica, step])
    x = rng.standard_normal(size=(batch, hw, hw, 3)).astype(np.float32)
    y = rng.integers(0, classes, size=(batch,)).astype(np.int32)
    return x, y


class SyncPSJob:
    def __init__(self, tf_config: str, gpu_rank: int = 0, n_gpus: int = 1,
                 cuda_device: Optional[int] = None, stream: int = 0):
        self.lib = L.load()
        self.h = C.c_void_p()
        self.tf_config = tf_config if isinstance(tf_config, str) else json.dumps(tf_config)
        rc = self.lib.job_create(self.tf_config.encode(), gpu_rank, n_gpus, C.byref(self.h))
        self.lib.check(rc)
        self.gpu_rank, self.n_gpus = gpu_rank, n_gpus
        self.batch = 0
        self._keep = []
        if cuda_device is not None:
            self.bind(cuda_device, stream)

    # -- lifecycle ---------------------------------------------------------------------
    def bind(self, cuda_device: int, stream: int = 0) -> None:
        self._ck(self.lib.job_bind(self.h, cuda_device, stream))

    def close(self) -> None:
        if self.h:
            self.lib.job_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc: int) -> None:
        self.lib.check(rc, self.h)

    # -- cluster -------------------------------------------------------------------------
    @property
    def n_replicas(self) -> int:
        return self.lib.num_replicas(self.h)

    @property
    def n_ps(self) -> int:
        return self.lib.num_ps(self.h)

    @property
    def local_replicas(self) -> List[int]:
        return [self.lib.local_replica(self.h, s) for s in range(self.lib.num_local_replicas(self.h))]

    def num_tasks(self, job: str) -> int:
        return self.lib.cluster_num_tasks(self.h, job.encode())

    # -- variables -------------------------------------------------------------------------
    def register(self, name: str, shape: Sequence[int], op_type: str = "VariableV2",
                 dtype: int = L.DT_F32, trainable: bool = True):
        arr = (C.c_int64 * max(1, len(shape)))(*shape)
        var, task = C.c_int(-1), C.c_int(-1)
        self._ck(self.lib.var_register(self.h, name.encode(), op_type.encode(), dtype, arr,
                                       len(shape), int(trainable), C.byref(var), C.byref(task)))
        return var.value, task.value

    def finalize(self, layout: str = "logical") -> None:
        self._ck(self.lib.vars_finalize(self.h, L.LAYOUTS[layout]))

    def placement(self, var: int) -> Dict[str, int]:
        t, g, off, cnt = C.c_int(), C.c_int(), C.c_int64(), C.c_int64()
        self._ck(self.lib.placement(self.h, var, C.byref(t), C.byref(g), C.byref(off), C.byref(cnt)))
        return {"ps_task": t.value, "gpu": g.value, "offset": off.value, "count": cnt.value}

    @property
    def n_vars(self) -> int:
        return self.lib.num_vars(self.h)

    @property
    def flat_count(self) -> int:
        return self.lib.flat_count(self.h)

    def set_var(self, var: int, value: np.ndarray) -> None:
        a = np.ascontiguousarray(value, dtype=np.float32)
        self._ck(self.lib.var_set(self.h, var, a.ctypes.data, a.nbytes))

    def _get(self, fn, var: int, *pre) -> np.ndarray:
        cnt = self.placement(var)["count"]
        out = np.empty(cnt, dtype=np.float32)
        self._ck(fn(self.h, *pre, var, out.ctypes.data, out.nbytes))
        return out

    def get_var(self, var: int) -> np.ndarray:
        return self._get(self.lib.var_get, var)

    def get_accum(self, var: int) -> np.ndarray:
        return self._get(self.lib.accum_get, var)

    def get_grad(self, slot: int, var: int) -> np.ndarray:
        return self._get(self.lib.grad_get, var, slot)

    def set_grad(self, slot: int, var: int, value: np.ndarray) -> None:
        a = np.ascontiguousarray(value, dtype=np.float32)
        self._ck(self.lib.grad_set(self.h, slot, var, a.ctypes.data, a.nbytes))

    def buffer(self, which: int, slot: int = 0):
        p, b = C.c_void_p(), C.c_size_t()
        self._ck(self.lib.job_buffer(self.h, which, slot, C.byref(p), C.byref(b)))
        return p.value, b.value

    def optimizer(self, kind: str = "sgd", lr: float = 0.01, momentum: float = 0.9) -> None:
        self._ck(self.lib.opt_config(self.h, L.OPT_KINDS[kind], lr, momentum))

    @property
    def global_step(self) -> int:
        return self.lib.global_step(self.h)

    @global_step.setter
    def global_step(self, step: int) -> None:
        self._ck(self.lib.global_step_set(self.h, int(step)))

    # -- raw flat-space access (checkpoints) -------------------------------------------------
    def owned_ranges(self):
        """[(offset, count)] of the flat-space ranges whose update this GPU rank applies."""
        n = C.c_int()
        self._ck(self.lib.owned_ranges(self.h, None, None, 0, C.byref(n)))
        offs, cnts = (C.c_int64 * max(1, n.value))(), (C.c_int64 * max(1, n.value))()
        self._ck(self.lib.owned_ranges(self.h, offs, cnts, n.value, C.byref(n)))
        return [(offs[i], cnts[i]) for i in range(n.value)]

    def flat_read(self, which: int, offset: int, count: int) -> np.ndarray:
        out = np.empty(count, dtype=np.float32)
        self._ck(self.lib.flat_read(self.h, which, offset, count, out.ctypes.data))
        return out

    def flat_write(self, which: int, offset: int, values: np.ndarray) -> None:
        a = np.ascontiguousarray(values, dtype=np.float32)
        self._ck(self.lib.flat_write(self.h, which, offset, a.size, a.ctypes.data))

    # -- model + step ----------------------------------------------------------------------
    def build_mnist(self, batch_per_replica: int = 128, math: str = "fp32",
                    layout: str = "logical", seed: Optional[int] = 0) -> None:
        self._ck(self.lib.model_mnist_cnn(self.h, batch_per_replica, L.MATH_MODES[math]))
        self.batch = batch_per_replica
        self.finalize(layout)
        if seed is not None:
            for v, w in enumerate(mnist_init(seed)):
                self.set_var(v, w)

    def var_info(self, var: int):
        name = C.create_string_buffer(256)
        shape = (C.c_int64 * 8)()
        rank, tr = C.c_int(), C.c_int()
        self._ck(self.lib.var_info(self.h, var, name, len(name), shape, C.byref(rank), C.byref(tr)))
        return name.value.decode(), tuple(shape[i] for i in range(rank.value)), bool(tr.value)

    def build_resnet50(self, batch_per_replica: int, image_hw: int = 224, classes: int = 1000,
                       layout: str = "balanced", seed: Optional[int] = 0) -> None:
        """BASELINE config 4.  Weights: He-normal convolutions, gamma 1 / beta 0, fc N(0, 0.01),
        all from numpy with a fixed seed (resnet50_init) and uploaded."""
        self._ck(self.lib.model_resnet50(self.h, batch_per_replica, image_hw, classes))
        self.batch, self.classes, self.image_hw = batch_per_replica, classes, image_hw
        self.finalize(layout)
        if seed is not None:
            for v, w in resnet50_init(self, seed).items():
                self.set_var(v, w)

    def tensor(self, name: str, slot: int = 0) -> np.ndarray:
        n = C.c_size_t()
        cap = 1 << 20
        while True:
            out = np.empty(cap, dtype=np.float32)
            rc = self.lib.tensor_get(self.h, slot, name.encode(), out.ctypes.data, cap, C.byref(n))
            if rc == 5 and n.value > cap:
                cap = n.value
                continue
            self._ck(rc)
            return out[:n.value]

    def forward_backward(self, slot: int, images_dev: int, labels_dev: int) -> None:
        self._ck(self.lib.forward_backward(self.h, slot, images_dev, labels_dev))

    def logits(self, slot: int = 0) -> np.ndarray:
        out = np.empty((self.batch, getattr(self, "classes", 10)), dtype=np.float32)
        self._ck(self.lib.logits_get(self.h, slot, out.ctypes.data, out.nbytes))
        return out

    def step(self, images_host: Sequence, labels_host: Sequence) -> float:
        """One synchronous step from HOST buffers (numpy arrays or pinned torch tensors),
        one pair per local replica slot; returns the mean loss over all replicas."""
        loss = C.c_float()
        xi, yi = L.ptr_array(images_host), L.ptr_array(labels_host)
        self._ck(self.lib.step_dense(self.h, xi, yi, C.byref(loss)))
        return loss.value

    def step_dev(self, images_dev: Sequence, labels_dev: Sequence, want_loss: bool = False):
        loss = C.c_float()
        xi, yi = L.ptr_array(images_dev), L.ptr_array(labels_dev)
        self._ck(self.lib.step_dense_dev(self.h, xi, yi, C.byref(loss) if want_loss else None))
        return loss.value if want_loss else None

    def push_apply_pull(self, with_barriers: bool = True) -> None:
        self._ck(self.lib.push_apply_pull(self.h, int(with_barriers)))

    def apply_all(self, n_summed: int) -> None:
        self._ck(self.lib.apply_all(self.h, n_summed))

    def grads_tensor(self, slot: int = 0):
        """The flat gradient buffer of a slot as a torch tensor view (no copy), for handing it to
        torch.distributed collectives in the NCCL baseline."""
        import torch
        ptr, nbytes = self.buffer(L.BUF_GRADS, slot)

        class _View:
            __cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4", "data": (ptr, False), "version": 3,
                                        "strides": None}
        return torch.as_tensor(_View(), device="cuda")

    def step_nccl(self, images_dev, labels_dev, dist) -> None:
        """Baseline exchange: forward/backward here, NCCL all-reduce of the flat gradient, then every
        rank applies the whole parameter vector (no peer-memory kernel, no CUDA graph)."""
        for s in range(len(images_dev)):
            self.forward_backward(s, L.ptr(images_dev[s]), L.ptr(labels_dev[s]))
        if not hasattr(self, "_gt"):
            self._gt = self.grads_tensor(0)
        if self.n_gpus > 1:
            dist.all_reduce(self._gt)
        self.apply_all(self.n_replicas)

    def apply_owned(self, n_summed: int) -> None:
        self._ck(self.lib.apply_owned(self.h, n_summed))

    def barrier(self) -> None:
        self._ck(self.lib.job_barrier(self.h))

    def loss(self, replica: int = -1) -> float:
        out = C.c_float()
        self._ck(self.lib.loss_get(self.h, replica, C.byref(out)))
        return out.value

    def sync(self) -> None:
        self._ck(self.lib.job_sync(self.h))

    def profile(self, on: bool = True) -> None:
        self._ck(self.lib.profile_enable(self.h, int(on)))

    def profile_report(self) -> Dict[str, Dict[str, float]]:
        buf = C.create_string_buffer(1 << 16)
        self._ck(self.lib.profile_report(self.h, buf, len(buf)))
        return json.loads(buf.value.decode())

    @property
    def launches(self) -> int:
        return self.lib.launch_count(self.h)

    # -- peers -----------------------------------------------------------------------------
    def ipc_export(self) -> bytes:
        buf = C.create_string_buffer(8 + 4 * L.MAX_REPLICAS + L.IPC_HANDLE_BYTES * (3 + L.MAX_REPLICAS))
        n = C.c_size_t()
        self._ck(self.lib.job_ipc_export(self.h, buf, len(buf), C.byref(n)))
        return buf.raw[:n.value]

    def ipc_open(self, peer_rank: int, blob: bytes) -> None:
        self._ck(self.lib.job_ipc_open(self.h, peer_rank, blob, len(blob)))

    def attach_local_peer(self, peer_rank: int, peer: "SyncPSJob") -> None:
        self._ck(self.lib.job_attach_local_peer(self.h, peer_rank, peer.h))

    def connect_peers(self, dist=None) -> None:
        """Exchange peer-memory handles over torch.distributed (any backend) and map every
        other rank's buffers."""
        if self.n_gpus == 1:
            return
        import torch.distributed as dist_mod
        dist = dist or dist_mod
        blobs: List[Optional[bytes]] = [None] * self.n_gpus
        dist.all_gather_object(blobs, self.ipc_export())
        for r, b in enumerate(blobs):
            if r != self.gpu_rank:
                self.ipc_open(r, b)
        dist.barrier()

    # -- smoke -----------------------------------------------------------------------------
    def smoke(self, all_tasks: bool = True) -> np.ndarray:
        n = C.c_int()
        out = np.zeros((4096, 100), dtype=np.int32)
        self._ck(self.lib.smoke(self.h, int(all_tasks), out.ctypes.data, out.size, C.byref(n)))
        return out[:n.value].reshape(n.value, 10, 10)
