"""Rendezvous of the replica processes of one TfJob on one box, over a Unix-domain socket.

The reference's replicas find each other through per-task Service DNS names
(pkg/trainer/replicas.go:157-186) and TensorFlow's gRPC servers.  On one box the GPU-rank
processes need exactly two things from each other before the first step: every rank's CUDA-IPC
handle blob (``tfg_job_ipc_export`` -> ``tfg_job_ipc_open``) and a barrier.  The launcher
(:class:`k8s_b200.backend.ProcessBackend`) runs a :class:`Server` whose socket path travels to the
replicas in ``TFGRAFT_RENDEZVOUS``; a :class:`Client` offers the two calls
:meth:`SyncPSJob.connect_peers` uses (``all_gather_object`` and ``barrier``, the subset of
``torch.distributed`` it needs), so no torchrun is involved.

Wire format: one JSON object per line.  ``{"op": "gather", "key": k, "rank": r, "world": n,
"data": base64}`` is answered, once all ``n`` ranks of key ``k`` have arrived, with ``{"data":
[base64 per rank]}``.  ``{"op": "wait_done"}`` is answered when the launcher calls
:meth:`Server.finish` (how a default-PS process learns that the job is over, the counterpart of
the pod deletion that ends ``server.join()`` in the reference).
"""
from __future__ import annotations

import base64
import json
import os
import pickle
import socket
import threading
import time
from typing import Any, Dict, List, Optional

ENV = "TFGRAFT_RENDEZVOUS"


class Server:
    def __init__(self, path: str):
        self.path = path
        if os.path.exists(path):
            os.unlink(path)
        self.sock = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        self.sock.bind(path)
        self.sock.listen(128)
        self.lock = threading.Condition()
        self.slots: Dict[str, Dict[int, str]] = {}
        self.done = False
        self.exit_code = 0
        self.closed = False
        self.thread = threading.Thread(target=self._accept, daemon=True)
        self.thread.start()

    def _accept(self) -> None:
        while not self.closed:
            try:
                conn, _ = self.sock.accept()
            except OSError:
                return
            threading.Thread(target=self._serve, args=(conn,), daemon=True).start()

    def _serve(self, conn: socket.socket) -> None:
        f = conn.makefile("rwb")
        try:
            for line in f:
                msg = json.loads(line)
                if msg["op"] == "gather":
                    key, world = msg["key"], int(msg["world"])
                    with self.lock:
                        self.slots.setdefault(key, {})[int(msg["rank"])] = msg["data"]
                        self.lock.notify_all()
                        deadline = time.time() + float(msg.get("timeout", 120.0))
                        while len(self.slots[key]) < world and not self.closed:
                            left = deadline - time.time()
                            if left <= 0:
                                break
                            self.lock.wait(left)
                        got = self.slots[key]
                        reply = ({"data": [got[r] for r in range(world)]} if len(got) >= world else
                                 {"error": "rendezvous %r: %d of %d ranks arrived" % (key, len(got), world)})
                elif msg["op"] == "wait_done":
                    with self.lock:
                        while not self.done and not self.closed:
                            self.lock.wait(1.0)
                        reply = {"exit_code": self.exit_code}
                else:
                    reply = {"error": "unknown op %r" % msg["op"]}
                f.write((json.dumps(reply) + "\n").encode())
                f.flush()
        except (OSError, ValueError):
            pass
        finally:
            try:
                conn.close()
            except OSError:
                pass

    def finish(self, exit_code: int = 0) -> None:
        with self.lock:
            self.done, self.exit_code = True, exit_code
            self.lock.notify_all()

    def close(self) -> None:
        self.closed = True
        with self.lock:
            self.lock.notify_all()
        try:
            self.sock.close()
        finally:
            if os.path.exists(self.path):
                os.unlink(self.path)


class Client:
    """The ``dist``-like object :meth:`SyncPSJob.connect_peers` takes."""

    def __init__(self, rank: int, world: int, path: Optional[str] = None, timeout: float = 120.0):
        self.rank, self.world, self.timeout = rank, world, timeout
        self.path = path or os.environ[ENV]
        self.seq = 0

    def _call(self, msg: Dict[str, Any]) -> Dict[str, Any]:
        deadline = time.time() + self.timeout
        while True:
            try:
                s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                s.connect(self.path)
                break
            except OSError:
                s.close()
                if time.time() > deadline:
                    raise
                time.sleep(0.05)
        with s:
            f = s.makefile("rwb")
            f.write((json.dumps(msg) + "\n").encode())
            f.flush()
            line = f.readline()
        if not line:
            raise ConnectionError("rendezvous server closed the connection")
        reply = json.loads(line)
        if "error" in reply:
            raise TimeoutError(reply["error"])
        return reply

    def all_gather_object(self, out: List[Any], obj: Any) -> None:
        self.seq += 1
        reply = self._call({"op": "gather", "key": "g%d" % self.seq, "rank": self.rank, "world": self.world,
                            "timeout": self.timeout, "data": base64.b64encode(pickle.dumps(obj)).decode()})
        for r, d in enumerate(reply["data"]):
            out[r] = pickle.loads(base64.b64decode(d))

    def barrier(self) -> None:
        self.all_gather_object([None] * self.world, None)

    def wait_done(self) -> int:
        s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        s.connect(self.path)
        with s:
            f = s.makefile("rwb")
            f.write((json.dumps({"op": "wait_done"}) + "\n").encode())
            f.flush()
            line = f.readline()
        return int(json.loads(line)["exit_code"]) if line else 0
