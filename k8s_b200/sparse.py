"""A partitioned embedding table on the GPU ranks of one box: the sparse PS path of a TfJob
(SURVEY.md par. 8a A15, par. 8e; north-star subsystem (7)).

[TF1.3] ``embedding_lookup`` over a partitioned variable: shard ``p = id mod S`` / local row
``id div S`` ("mod"), or contiguous ranges ("div").  Shard k is the variable piece placed on
``/job:ps/task:k``, i.e. owned by GPU ``k % n_gpus`` (the same rule as the dense shards,
``tfg_ps_gpu``).  Every rank hosts one training replica's batch; all arithmetic is behind the C ABI
(``tfg_sparse_gather_multi`` / ``tfg_sparse_scatter_sub_multi``), torch only owns the device buffers.

  * pull   -- ``lookup(ids)``: one launch per worker reads every row straight from its owner's shard
             (peer loads), no all-to-all;
  * push   -- ``apply_gradients(rows, lr)``: each shard owner applies ``scatter_sub`` over the
             concatenation of all replicas' (ids, rows) in (replica, position) order, reading them in
             place from the replicas' buffers (peer loads): bit-identical to the sequential loop.
Cross-rank ordering uses the job's stream-ordered barrier (``tfg_job_barrier``).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import numpy as np

from . import lib as L


class ShardedEmbedding:
    def __init__(self, job, vocab: int, dim: int, batch: int, strategy: str = "mod", n_shards: Optional[int] = None,
                 stream=None, chunk_long_runs: bool = False):
        """chunk_long_runs: give the update the extra scratch with which a row hit by more than 64 ids
        (heavy-tailed id distributions) is summed in blocks of 128 ids by many warps instead of one id
        after another by one warp -- deterministic and as accurate, but not bit-identical to the
        sequential loop (include/tfgraft.h, tfg_sparse_scratch_bytes_fast)."""
        import torch
        self.torch = torch
        self.job, self.lib = job, job.lib
        self.vocab, self.dim, self.batch = vocab, dim, batch
        self.strategy = {"mod": L.PART_MOD, "div": L.PART_DIV}[strategy]
        self.n_shards = n_shards or max(1, job.n_ps)
        self.n_gpus, self.rank = job.n_gpus, job.gpu_rank
        self.stream = stream                       # a torch.cuda.Stream the job was bound to, or None (default stream)
        self.owned = [k for k in range(self.n_shards) if k % self.n_gpus == self.rank]
        # buffers that other ranks map are allocations of their own (tfg_dev_alloc), not pieces of torch's
        # caching allocator: a CUDA-IPC handle names a whole allocation
        self._allocs: List[int] = []
        self.shards = {k: self._shared(max(1, self.lib.sparse_shard_rows(vocab, self.n_shards, k, self.strategy)) * dim, "<f4")
                       for k in self.owned}
        self.ids = self._shared(batch, "<i8")
        self.rows = self._shared(batch * dim, "<f4")
        self.out = torch.zeros(batch * dim, device="cuda")
        n_total = batch * self.n_gpus
        need = self.lib.sparse_scratch_bytes_fast(n_total, dim) if chunk_long_runs else self.lib.sparse_scratch_bytes(n_total)
        self.scratch = torch.empty(need + 256, dtype=torch.uint8, device="cuda")
        self.bad = torch.zeros(1, dtype=torch.int32, device="cuda")
        self._opened: List[int] = []
        # pointer tables; filled by connect()
        self.shard_ptrs = [0] * self.n_shards
        self.ids_ptrs = [0] * self.n_gpus
        self.rows_ptrs = [0] * self.n_gpus
        for k in self.owned:
            self.shard_ptrs[k] = self.shards[k].data_ptr()
        self.ids_ptrs[self.rank], self.rows_ptrs[self.rank] = self.ids.data_ptr(), self.rows.data_ptr()

    def _shared(self, n: int, typestr: str):
        p = C.c_void_p()
        self.lib.check(self.lib.dev_alloc(n * int(typestr[2:]), C.byref(p)))
        self._allocs.append(p.value)
        addr = p.value

        class _View:
            __cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (addr, False), "version": 3, "strides": None}
        return self.torch.as_tensor(_View(), device="cuda")

    # -- peers -------------------------------------------------------------------------------
    def _export(self, t) -> bytes:
        buf = C.create_string_buffer(L.IPC_HANDLE_BYTES)
        self.lib.check(self.lib.ipc_export(t.data_ptr(), buf))
        return buf.raw

    def _open(self, handle: bytes) -> int:
        p = C.c_void_p()
        self.lib.check(self.lib.ipc_open(handle, C.byref(p)))
        self._opened.append(p.value)
        return p.value

    def connect(self, dist) -> None:
        """Exchange CUDA-IPC handles of the shards and of the id / row staging buffers (any object
        with all_gather_object + barrier: torch.distributed or k8s_b200.rendezvous.Client)."""
        if self.n_gpus == 1:
            return
        mine = {"shards": {k: self._export(self.shards[k]) for k in self.owned},
                "ids": self._export(self.ids), "rows": self._export(self.rows)}
        everyone = [None] * self.n_gpus
        dist.all_gather_object(everyone, mine)
        for r, blob in enumerate(everyone):
            if r == self.rank:
                continue
            for k, h in blob["shards"].items():
                self.shard_ptrs[k] = self._open(h)
            self.ids_ptrs[r] = self._open(blob["ids"])
            self.rows_ptrs[r] = self._open(blob["rows"])
        dist.barrier()

    def attach_local(self, others: List["ShardedEmbedding"]) -> None:
        """Single-process stand-in used by the 1-GPU tests: the other ranks' tables live in this
        process on the same device, so their pointers are usable as they are."""
        for o in others:
            for k in o.owned:
                self.shard_ptrs[k] = o.shards[k].data_ptr()
            self.ids_ptrs[o.rank], self.rows_ptrs[o.rank] = o.ids.data_ptr(), o.rows.data_ptr()

    def _s(self) -> int:
        return self.stream.cuda_stream if self.stream is not None else 0

    # -- table contents (tests, checkpoints) ----------------------------------------------------
    def set_shard(self, k: int, values: np.ndarray) -> None:
        self.shards[k].copy_(self.torch.from_numpy(np.ascontiguousarray(values, dtype=np.float32).ravel()))

    def get_shard(self, k: int) -> np.ndarray:
        return self.shards[k].cpu().numpy().reshape(-1, self.dim)

    # -- the two halves of a sparse step ---------------------------------------------------------
    def stage_ids(self, ids: np.ndarray) -> None:
        self.ids.copy_(self.torch.from_numpy(np.ascontiguousarray(ids, dtype=np.int64)))

    def lookup(self) -> "np.ndarray":
        """Rows of the staged ids, [batch, dim] (device tensor view)."""
        self.bad.zero_()
        self.lib.check(self.lib.sparse_gather_multi(L.ptr_array(self.shard_ptrs), self.vocab, self.dim, self.n_shards,
                                                    self.strategy, self.ids.data_ptr(), self.batch, self.out.data_ptr(),
                                                    self.bad.data_ptr(), self._s()))
        return self.out.view(self.batch, self.dim)

    def stage_rows(self, rows: np.ndarray) -> None:
        self.rows.copy_(self.torch.from_numpy(np.ascontiguousarray(rows, dtype=np.float32).ravel()))

    def apply_owned(self, scale: float) -> None:
        """scatter_sub of every replica's staged (ids, rows) into the shards this rank owns; the caller
        orders it after all ranks have staged (barrier) and before anyone restages."""
        for k in self.owned:
            self.lib.check(self.lib.sparse_scatter_sub_multi(
                self.shards[k].data_ptr(), self.vocab, self.dim, self.n_shards, k, self.strategy,
                L.ptr_array(self.ids_ptrs), L.ptr_array(self.rows_ptrs), self.n_gpus, self.batch, scale,
                (self.scratch.data_ptr() + 255) // 256 * 256, self.scratch.numel() - 256, self._s()))

    def apply_gradients(self, rows: np.ndarray, lr: float) -> None:
        """One synchronous sparse update across the ranks: stage, barrier, owners apply, barrier."""
        self.stage_rows(rows)
        if self.n_gpus > 1:
            self.job.barrier()
        self.apply_owned(lr)
        if self.n_gpus > 1:
            self.job.barrier()

    def close(self) -> None:
        """Unmap the peers' buffers and free the shared ones (call on every rank after a barrier: nobody may
        still be reading them)."""
        for p in self._opened:
            self.lib.ipc_close(p)
        self._opened = []
        self.torch.cuda.synchronize()
        self.shards, self.ids, self.rows = {}, None, None
        for p in self._allocs:
            self.lib.dev_free(p)
        self._allocs = []
