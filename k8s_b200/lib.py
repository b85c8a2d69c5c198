"""ctypes binding of libtfgraft (include/tfgraft.h) -- the only way Python reaches the
CUDA kernels.  There is no CPU fallback: a missing library raises at load time, and every
compute entry point fails loudly without a usable CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os
import re
from typing import List, Optional, Sequence

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libtfgraft.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "tfgraft.h")

# enums of include/tfgraft.h
OPT_SGD, OPT_MOMENTUM, OPT_NESTEROV = 0, 1, 2
LAYOUT_LOGICAL, LAYOUT_BALANCED = 0, 1
MATH_FP32, MATH_BF16, MATH_TF32, MATH_TF32X3 = 0, 1, 2, 3
PART_MOD, PART_DIV = 0, 1
DT_F32, DT_I64 = 0, 1
BUF_PARAMS, BUF_GRADS, BUF_ACCUM, BUF_LOSS, BUF_FLAGS = 0, 1, 2, 3, 4
MAX_REPLICAS, MAX_GPUS, IPC_HANDLE_BYTES = 64, 16, 64
SPARSE_STATIC_MAX_IDS, SPARSE_STATIC_SCRATCH_BYTES = 65536, 1204480
OPT_KINDS = {"sgd": OPT_SGD, "momentum": OPT_MOMENTUM, "nesterov": OPT_NESTEROV}
MATH_MODES = {"fp32": MATH_FP32, "bf16": MATH_BF16, "tf32": MATH_TF32, "tf32x3": MATH_TF32X3}
LAYOUTS = {"logical": LAYOUT_LOGICAL, "balanced": LAYOUT_BALANCED}


class TfgError(RuntimeError):
    """A non-zero return of the C ABI.  ``code`` follows the reference's exit-code
    contract (pkg/trainer/training.go:203-238): 1..127 permanent, >=128 retryable."""

    def __init__(self, code: int, message: str):
        super().__init__("libtfgraft rc=%d: %s" % (code, message))
        self.code = code
        self.message = message

    @property
    def retryable(self) -> bool:
        return self.code >= 128


def declared_symbols(header: str = HEADER_PATH) -> List[str]:
    """Every function name include/tfgraft.h declares."""
    text = open(header).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tfg_[a-z0-9_]+)\s*\(", text)))


_p = C.c_void_p
_SIGS = {
    "tfg_version": (C.c_int, []),
    "tfg_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "tfg_last_error": (C.c_char_p, [_p]),
    "tfg_job_create": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.POINTER(_p)]),
    "tfg_job_destroy": (C.c_int, [_p]),
    "tfg_job_bind": (C.c_int, [_p, C.c_int, _p]),
    "tfg_cluster_num_tasks": (C.c_int, [_p, C.c_char_p]),
    "tfg_num_replicas": (C.c_int, [_p]),
    "tfg_num_ps": (C.c_int, [_p]),
    "tfg_replica_gpu": (C.c_int, [_p, C.c_int]),
    "tfg_ps_gpu": (C.c_int, [_p, C.c_int]),
    "tfg_num_local_replicas": (C.c_int, [_p]),
    "tfg_local_replica": (C.c_int, [_p, C.c_int]),
    "tfg_parse_cluster_spec": (C.c_int, [C.c_char_p, C.c_char_p, C.c_size_t]),
    "tfg_var_register": (C.c_int, [_p, C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int64),
                                   C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "tfg_num_vars": (C.c_int, [_p]),
    "tfg_var_info": (C.c_int, [_p, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_int64), C.POINTER(C.c_int),
                               C.POINTER(C.c_int)]),
    "tfg_model_resnet50": (C.c_int, [_p, C.c_int, C.c_int, C.c_int]),
    "tfg_tensor_get": (C.c_int, [_p, C.c_int, C.c_char_p, _p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "tfg_vars_finalize": (C.c_int, [_p, C.c_int]),
    "tfg_placement": (C.c_int, [_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "tfg_flat_count": (C.c_int64, [_p]),
    "tfg_var_set": (C.c_int, [_p, C.c_int, _p, C.c_size_t]),
    "tfg_var_get": (C.c_int, [_p, C.c_int, _p, C.c_size_t]),
    "tfg_accum_get": (C.c_int, [_p, C.c_int, _p, C.c_size_t]),
    "tfg_grad_set": (C.c_int, [_p, C.c_int, C.c_int, _p, C.c_size_t]),
    "tfg_grad_get": (C.c_int, [_p, C.c_int, C.c_int, _p, C.c_size_t]),
    "tfg_job_buffer": (C.c_int, [_p, C.c_int, C.c_int, C.POINTER(_p), C.POINTER(C.c_size_t)]),
    "tfg_opt_config": (C.c_int, [_p, C.c_int, C.c_float, C.c_float]),
    "tfg_global_step": (C.c_int64, [_p]),
    "tfg_global_step_set": (C.c_int, [_p, C.c_int64]),
    "tfg_owned_ranges": (C.c_int, [_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_int, C.POINTER(C.c_int)]),
    "tfg_flat_read": (C.c_int, [_p, C.c_int, C.c_int64, C.c_int64, _p]),
    "tfg_flat_write": (C.c_int, [_p, C.c_int, C.c_int64, C.c_int64, _p]),
    "tfg_job_ipc_export": (C.c_int, [_p, _p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "tfg_job_ipc_open": (C.c_int, [_p, C.c_int, _p, C.c_size_t]),
    "tfg_job_attach_local_peer": (C.c_int, [_p, C.c_int, _p]),
    "tfg_job_barrier": (C.c_int, [_p]),
    "tfg_push_apply_pull": (C.c_int, [_p, C.c_int]),
    "tfg_apply_owned": (C.c_int, [_p, C.c_int]),
    "tfg_apply_all": (C.c_int, [_p, C.c_int]),
    "tfg_dense_apply": (C.c_int, [C.c_int, _p, _p, C.POINTER(_p), C.c_int, C.c_float, C.c_int64,
                                  C.c_float, C.c_float, C.POINTER(_p), C.c_int, _p]),
    "tfg_sparse_shard_rows": (C.c_int64, [C.c_int64, C.c_int, C.c_int, C.c_int]),
    "tfg_sparse_gather": (C.c_int, [_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, _p,
                                    C.c_int64, _p, _p, _p]),
    "tfg_sparse_scratch_bytes": (C.c_size_t, [C.c_int64]),
    "tfg_sparse_scratch_bytes_fast": (C.c_size_t, [C.c_int64, C.c_int]),
    "tfg_sparse_scatter_sub": (C.c_int, [_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, _p,
                                         C.c_int64, _p, C.c_float, _p, C.c_size_t, _p]),
    "tfg_sparse_gather_multi": (C.c_int, [C.POINTER(_p), C.c_int64, C.c_int, C.c_int, C.c_int, _p, C.c_int64, _p, _p, _p]),
    "tfg_sparse_scatter_sub_multi": (C.c_int, [_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_p),
                                               C.POINTER(_p), C.c_int, C.c_int64, C.c_float, _p, C.c_size_t, _p]),
    "tfg_dev_alloc": (C.c_int, [C.c_size_t, C.POINTER(_p)]),
    "tfg_dev_free": (C.c_int, [_p]),
    "tfg_ipc_export": (C.c_int, [_p, _p]),
    "tfg_ipc_open": (C.c_int, [_p, C.POINTER(_p)]),
    "tfg_ipc_close": (C.c_int, [_p]),
    "tfg_sparse_scatter_sub_static": (C.c_int, [_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, _p,
                                                C.c_int64, _p, C.c_float, _p]),
    "tfg_model_mnist_cnn": (C.c_int, [_p, C.c_int, C.c_int]),
    "tfg_forward_backward": (C.c_int, [_p, C.c_int, _p, _p]),
    "tfg_logits_get": (C.c_int, [_p, C.c_int, _p, C.c_size_t]),
    "tfg_step_dense": (C.c_int, [_p, C.POINTER(_p), C.POINTER(_p), C.POINTER(C.c_float)]),
    "tfg_step_dense_dev": (C.c_int, [_p, C.POINTER(_p), C.POINTER(_p), C.POINTER(C.c_float)]),
    "tfg_loss_get": (C.c_int, [_p, C.c_int, C.POINTER(C.c_float)]),
    "tfg_job_sync": (C.c_int, [_p]),
    "tfg_launch_count": (C.c_int64, [_p]),
    "tfg_gemm_tc": (C.c_int, [C.c_int, _p, C.c_int, _p, C.c_int, _p, C.c_int, C.c_int, C.c_int, C.c_int64,
                              C.c_int64, C.c_int, C.c_int, _p, _p, C.c_int, _p]),
    "tfg_gemm_tc_x3": (C.c_int, [_p, _p, C.c_int, _p, _p, C.c_int, _p, C.c_int, C.c_int, C.c_int, C.c_int64,
                                 C.c_int64, C.c_int, C.c_int, C.c_int, _p, _p, C.c_int, _p]),
    "tfg_tf32_split_lo": (C.c_int, [_p, _p, C.c_int64, _p]),
    "tfg_profile_enable": (C.c_int, [_p, C.c_int]),
    "tfg_profile_report": (C.c_int, [_p, C.c_char_p, C.c_size_t]),
    "tfg_smoke": (C.c_int, [_p, C.c_int, _p, C.c_size_t, C.POINTER(C.c_int)]),
}


class Lib:
    """The loaded shared library with typed entry points."""

    def __init__(self, path: str = LIB_PATH):
        if not os.path.exists(path):
            raise ImportError(
                "libtfgraft.so is not built (%s). Run `make -C k8s_b200/csrc` or "
                "`python -c 'import __graft_entry__ as g; g.build()'`; there is no CPU "
                "fallback for the CUDA path." % path)
        self.path = path
        self.dll = C.CDLL(path)
        for name, (res, args) in _SIGS.items():
            fn = getattr(self.dll, name)  # AttributeError if the header and the .so disagree
            fn.restype = res
            fn.argtypes = args
            setattr(self, name[4:], fn)

    def error(self, job=None) -> str:
        s = self.last_error(job)
        return s.decode("utf-8", "replace") if s else ""

    def check(self, rc: int, job=None) -> None:
        if rc != 0:
            raise TfgError(rc, self.error(job) or self.error(None))

    def device_count_or_zero(self) -> int:
        n = C.c_int(0)
        rc = self.device_count(C.byref(n))
        return n.value if rc == 0 else 0

    def device_count_checked(self) -> int:
        n = C.c_int(0)
        self.check(self.device_count(C.byref(n)))
        return n.value

    def parse_cluster_spec_json(self, spec: str) -> str:
        buf = C.create_string_buffer(1 << 16)
        rc = self.parse_cluster_spec(spec.encode(), buf, len(buf))
        if rc != 0:
            raise ValueError(self.error(None))
        return buf.value.decode()


_LIB: Optional[Lib] = None


def load() -> Lib:
    global _LIB
    if _LIB is None:
        _LIB = Lib()
    return _LIB


def ptr(t) -> int:
    """Raw address of a torch tensor / numpy array / int."""
    if t is None:
        return 0
    if isinstance(t, int):
        return t
    if hasattr(t, "data_ptr"):
        return t.data_ptr()
    return t.ctypes.data


def ptr_array(items: Sequence) -> "C.Array":
    arr = (C.c_void_p * max(1, len(items)))()
    for i, it in enumerate(items):
        arr[i] = ptr(it)
    return arr
