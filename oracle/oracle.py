"""ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy-facing loader of oracle/liboracle.so (the plain-C restatement in oracle/oracle.c).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this.  PARITY UNPINNED except for ``smoke`` -- see the header of oracle.c.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Sequence

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "liboracle.so")

SGD, MOMENTUM, NESTEROV = 0, 1, 2
MOD, DIV = 0, 1
KINDS = {"sgd": SGD, "momentum": MOMENTUM, "nesterov": NESTEROV}


def build() -> None:
    subprocess.check_call(["make", "-C", HERE, "-s"])


_dll = None


def dll():
    global _dll
    if _dll is None:
        if not os.path.exists(SO):
            build()
        _dll = C.CDLL(SO)
        _dll.orc_shard_rows.restype = C.c_int64
        _dll.orc_shard_rows.argtypes = [C.c_int64, C.c_int, C.c_int, C.c_int]
        _dll.orc_sparse_gather.restype = C.c_int64
    return _dll


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a


def smoke(n_tasks: int) -> np.ndarray:
    out = np.zeros((n_tasks, 100), dtype=np.int32)
    dll().orc_smoke(out.ctypes.data_as(C.c_void_p), C.c_int(n_tasks))
    return out.reshape(n_tasks, 10, 10)


def round_robin(op_types: Sequence[str], n_ps: int) -> List[int]:
    flags = np.array([t in ("Variable", "VariableV2", "VarHandleOp") for t in op_types], dtype=np.int32)
    out = np.zeros(len(op_types), dtype=np.int32)
    dll().orc_round_robin(flags.ctypes.data_as(C.c_void_p), C.c_int(len(op_types)), C.c_int(n_ps),
                          out.ctypes.data_as(C.c_void_p))
    return out.tolist()


def dense_apply(kind: int, var: np.ndarray, accum, grads: Sequence[np.ndarray], divisor: float,
                lr: float, mu: float):
    """In place on ``var`` (and ``accum``); grads summed in list order."""
    assert var.dtype == np.float32 and var.flags.c_contiguous
    gs = [_f32(g) for g in grads]
    arr = (C.c_void_p * len(gs))(*[g.ctypes.data for g in gs])
    ap = accum.ctypes.data_as(C.c_void_p) if accum is not None else None
    dll().orc_dense_apply(C.c_int(kind), var.ctypes.data_as(C.c_void_p), ap, arr, C.c_int(len(gs)),
                          C.c_float(divisor), C.c_int64(var.size), C.c_float(lr), C.c_float(mu))
    return var, accum


def locate(vocab: int, n_shards: int, strategy: int, id_: int):
    sh, row = C.c_int(), C.c_int64()
    dll().orc_locate(C.c_int64(vocab), C.c_int(n_shards), C.c_int(strategy), C.c_int64(id_),
                     C.byref(sh), C.byref(row))
    return sh.value, row.value


def shard_rows(vocab: int, n_shards: int, shard: int, strategy: int) -> int:
    return dll().orc_shard_rows(vocab, n_shards, shard, strategy)


def sparse_gather(table: np.ndarray, vocab: int, n_shards: int, shard: int, strategy: int,
                  ids: np.ndarray, out: np.ndarray) -> int:
    assert table.dtype == np.float32 and out.dtype == np.float32 and ids.dtype == np.int64
    dim = table.shape[1]
    return dll().orc_sparse_gather(table.ctypes.data_as(C.c_void_p), C.c_int64(vocab), C.c_int(dim),
                                   C.c_int(n_shards), C.c_int(shard), C.c_int(strategy),
                                   ids.ctypes.data_as(C.c_void_p), C.c_int64(ids.size),
                                   out.ctypes.data_as(C.c_void_p))


def sparse_scatter_sub(table: np.ndarray, vocab: int, n_shards: int, shard: int, strategy: int,
                       ids: np.ndarray, rows: np.ndarray, scale: float) -> None:
    assert table.dtype == np.float32 and rows.dtype == np.float32 and ids.dtype == np.int64
    dim = table.shape[1]
    dll().orc_sparse_scatter_sub(table.ctypes.data_as(C.c_void_p), C.c_int64(vocab), C.c_int(dim),
                                 C.c_int(n_shards), C.c_int(shard), C.c_int(strategy),
                                 ids.ctypes.data_as(C.c_void_p), C.c_int64(ids.size),
                                 rows.ctypes.data_as(C.c_void_p), C.c_float(scale))
