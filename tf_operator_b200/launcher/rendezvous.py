"""Launcher-hosted rendezvous store and handle broker.

Replaces the DNS + TCPStore bootstrap of the reference stack (MASTER_ADDR:MASTER_PORT,
pkg/controller.v1/pytorch/pytorch.go:41-65): the replicas of a job meet here to publish the CUDA
IPC handles of their symmetric buffers (64-byte blobs), to pass the multicast (NVLS) file
descriptor from the rank that created it to the others, and to barrier. Values are bytes; ``get``
blocks until the key exists.

Life cycle (SURVEY.md 7.3 "hard part"; pkg/controller.v1/tensorflow/pod.go:140-159 re-creates a
retryable replica under the SAME name, py/kubeflow/tf_operator/simple_tfjob_tests.py:26-90
re-creates a whole job under the same name):

* every key lives under a *scope* ``<namespace>/<job>/<job uid>/<tag>`` - a re-created job has a
  new uid, so it can never read what its predecessor published;
* inside a scope, replicas ``join`` a *generation*. A rank that joins while it is already a
  member is a restarted process: it opens generation g+1 (and the keys and descriptors of
  generation g are dropped). Its peers are still running against generation g; their collectives
  time out, they exit with a retryable code, the operator re-creates them, and they join g+1.
  A waiter whose generation is superseded gets ``superseded`` and joins again;
* ``delete_prefix`` (called by the cluster when the job object is deleted) drops keys, counters,
  scopes and closes brokered descriptors.

Wire protocol (TCP and the Unix socket alike): one JSON object per line, values base64.
  {"op":"set","key":k,"value":b64} -> {"ok":true}
  {"op":"get","key":k,"timeout":s[,"scope":s,"gen":g]} -> {"ok":true,"value":b64} |
                                      {"ok":false,"error":"timeout"|"superseded"}
  {"op":"add","key":k,"delta":n}   -> {"ok":true,"value":n}
  {"op":"join","scope":s,"rank":r,"world":w,"timeout":t} -> {"ok":true,"value":generation}
  {"op":"delete_prefix","prefix":p}-> {"ok":true,"value":count}
Unix socket only (SCM_RIGHTS carries the descriptor next to the line):
  {"op":"fd_put","key":k} + fd     -> {"ok":true}
  {"op":"fd_get","key":k,"timeout":s} -> {"ok":true} + fd
"""
from __future__ import annotations

import base64
import json
import os
import socket
import socketserver
import tempfile
import threading
import time
import uuid
from typing import Dict, List, Optional


class Superseded(RuntimeError):
    """The generation this replica was joining has been replaced by a newer one."""


class Replaced(RuntimeError):
    """A newer incarnation of this very rank has joined: this process is the stale one."""


class _Scope:
    def __init__(self):
        self.gen = 0
        self.members: Dict[int, str] = {}     # rank -> identity of the process that joined


class RendezvousServer:
    def __init__(self, host: str = "127.0.0.1", port: int = 0, uds: bool = True):
        self._data: Dict[str, bytes] = {}
        self._counters: Dict[str, int] = {}
        self._fds: Dict[str, int] = {}
        self._scopes: Dict[str, _Scope] = {}
        self._cv = threading.Condition()
        outer = self

        class Handler(socketserver.StreamRequestHandler):
            def handle(self):
                for line in self.rfile:
                    try:
                        req = json.loads(line)
                        resp = outer._dispatch(req)
                    except Exception as e:  # noqa: BLE001
                        resp = {"ok": False, "error": repr(e)}
                    try:
                        self.wfile.write(json.dumps(resp).encode() + b"\n")
                        self.wfile.flush()
                    except OSError:
                        return

        class Server(socketserver.ThreadingTCPServer):
            allow_reuse_address = True
            daemon_threads = True

        self._srv = Server((host, port), Handler)
        self.host, self.port = self._srv.server_address[:2]
        self._thread: Optional[threading.Thread] = None
        # descriptor broker: a Unix stream socket next to the TCP one
        self.uds_path = ""
        self._uds: Optional[socket.socket] = None
        self._uds_dir = ""
        if uds:
            self._uds_dir = tempfile.mkdtemp(prefix="tfop-rdv-")
            self.uds_path = os.path.join(self._uds_dir, "broker.sock")
            self._uds = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            self._uds.bind(self.uds_path)
            self._uds.listen(64)
        self._stopping = False

    @property
    def addr(self) -> str:
        return f"{self.host}:{self.port}"

    # ------------------------------------------------------------------ operations
    def _wait_key(self, key: str, timeout: float, store, scope: str = "", gen: int = -1) -> str:
        """"" when the key is there, "timeout", or "superseded": the waiter named the rendezvous
        generation it belongs to (scope, gen) and that generation is over - a peer restarted, its
        keys are gone and nobody will ever publish the one waited for."""
        end = time.time() + timeout
        while key not in store:
            if scope:
                st = self._scopes.get(scope)
                if st is None or st.gen != gen:
                    return "superseded"
            left = end - time.time()
            if left <= 0:
                return "timeout"
            self._cv.wait(left)
        return ""

    def _dispatch(self, req):
        op = req.get("op")
        if op == "set":
            with self._cv:
                self._data[req["key"]] = base64.b64decode(req["value"])
                self._cv.notify_all()
            return {"ok": True}
        if op == "get":
            with self._cv:
                err = self._wait_key(req["key"], float(req.get("timeout", 60)), self._data,
                                     str(req.get("scope", "")), int(req.get("gen", -1)))
                if err:
                    return {"ok": False, "error": err}
                return {"ok": True, "value": base64.b64encode(self._data[req["key"]]).decode()}
        if op == "add":
            with self._cv:
                v = self._counters.get(req["key"], 0) + int(req.get("delta", 1))
                self._counters[req["key"]] = v
                self._data[req["key"]] = str(v).encode()
                self._cv.notify_all()
            return {"ok": True, "value": v}
        if op == "join":
            return self._join(req["scope"], int(req["rank"]), int(req["world"]),
                              float(req.get("timeout", 60)), str(req.get("ident", "")))
        if op == "delete_prefix":
            return {"ok": True, "value": self.delete_prefix(req["prefix"])}
        return {"ok": False, "error": f"unknown op {op!r}"}

    def _join(self, scope: str, rank: int, world: int, timeout: float, ident: str = ""):
        end = time.time() + timeout
        with self._cv:
            st = self._scopes.setdefault(scope, _Scope())
            if rank in st.members and not (ident and st.members[rank] == ident):
                # this rank was here before under another identity: the process restarted.
                # Everything the previous generation published (IPC handles of freed memory,
                # barrier counts) is stale.
                old = st.gen
                st.gen += 1
                st.members = {}
                self._drop_locked(f"{scope}/g{old}/")
            st.members[rank] = ident
            mine = st.gen
            self._cv.notify_all()
            while True:
                if st.gen != mine:
                    # superseded by a newer incarnation of my own rank: I am the zombie
                    if rank in st.members and st.members[rank] != ident:
                        return {"ok": False, "error": "replaced"}
                    return {"ok": False, "error": "superseded"}
                if len(st.members) >= world:
                    return {"ok": True, "value": mine}
                left = end - time.time()
                if left <= 0:
                    return {"ok": False, "error": "timeout"}
                self._cv.wait(left)

    def _drop_locked(self, prefix: str) -> int:
        ks = [k for k in self._data if k.startswith(prefix)]
        for k in ks:
            self._data.pop(k, None)
        for k in [k for k in self._counters if k.startswith(prefix)]:
            self._counters.pop(k, None)
        for k in [k for k in self._fds if k.startswith(prefix)]:
            try:
                os.close(self._fds.pop(k))
            except OSError:
                pass
            ks.append(k)
        return len(ks)

    def delete_prefix(self, prefix: str) -> int:
        """Drop every key, counter, scope and brokered descriptor under ``prefix``. The cluster
        calls this with ``<ns>/<job>/<uid>/`` when a job object is deleted."""
        with self._cv:
            n = self._drop_locked(prefix)
            for s in [s for s in self._scopes if s.startswith(prefix)]:
                del self._scopes[s]
            self._cv.notify_all()
        return n

    def keys(self, prefix: str = "") -> List[str]:
        with self._cv:
            return sorted(k for k in list(self._data) + list(self._fds) if k.startswith(prefix))

    # ------------------------------------------------------------------ descriptor broker
    def _uds_loop(self):
        while not self._stopping:
            try:
                conn, _ = self._uds.accept()
            except OSError:
                return
            threading.Thread(target=self._uds_conn, args=(conn,), daemon=True).start()

    def _uds_conn(self, conn: socket.socket):
        with conn:
            while True:
                try:
                    msg, fds, _, _ = socket.recv_fds(conn, 65536, 4)
                except OSError:
                    return
                if not msg:
                    for fd in fds:
                        os.close(fd)
                    return
                try:
                    req = json.loads(msg)
                    op = req.get("op")
                    if op == "fd_put":
                        if not fds:
                            raise ValueError("fd_put without a descriptor")
                        with self._cv:
                            old = self._fds.pop(req["key"], None)
                            if old is not None:
                                os.close(old)
                            self._fds[req["key"]] = fds[0]
                            for extra in fds[1:]:
                                os.close(extra)
                            self._cv.notify_all()
                        conn.sendall(b'{"ok": true}\n')
                    elif op == "fd_get":
                        with self._cv:
                            err = self._wait_key(req["key"], float(req.get("timeout", 60)), self._fds,
                                                 str(req.get("scope", "")), int(req.get("gen", -1)))
                            fd = os.dup(self._fds[req["key"]]) if not err else -1
                        if not err:
                            try:
                                socket.send_fds(conn, [b'{"ok": true}\n'], [fd])
                            finally:
                                os.close(fd)
                        else:
                            conn.sendall((json.dumps({"ok": False, "error": err}) + "\n").encode())
                    else:
                        for fd in fds:
                            os.close(fd)
                        conn.sendall((json.dumps(self._dispatch(req)) + "\n").encode())
                except Exception as e:  # noqa: BLE001
                    try:
                        conn.sendall((json.dumps({"ok": False, "error": repr(e)}) + "\n").encode())
                    except OSError:
                        return

    # ------------------------------------------------------------------ life cycle
    def start(self) -> "RendezvousServer":
        self._thread = threading.Thread(target=self._srv.serve_forever, daemon=True,
                                        name="tfop-rendezvous")
        self._thread.start()
        if self._uds is not None:
            threading.Thread(target=self._uds_loop, daemon=True, name="tfop-rendezvous-fd").start()
        return self

    def stop(self) -> None:
        self._stopping = True
        self._srv.shutdown()
        self._srv.server_close()
        if self._uds is not None:
            try:
                self._uds.close()
                os.unlink(self.uds_path)
                os.rmdir(self._uds_dir)
            except OSError:
                pass
        with self._cv:
            for fd in self._fds.values():
                try:
                    os.close(fd)
                except OSError:
                    pass
            self._fds.clear()


class RendezvousClient:
    def __init__(self, addr: str, timeout: float = 120.0, uds_path: str = ""):
        host, _, port = addr.rpartition(":")
        self._sock = socket.create_connection((host, int(port)), timeout=timeout + 30)
        self._f = self._sock.makefile("rwb")
        self._lock = threading.Lock()
        self.timeout = timeout
        self.uds_path = uds_path or os.environ.get("TFOP_RENDEZVOUS_UDS", "")
        self._uuid = uuid.uuid4().hex            # one identity per client = per replica process

    def _call(self, req):
        with self._lock:
            self._f.write(json.dumps(req).encode() + b"\n")
            self._f.flush()
            line = self._f.readline()
        if not line:
            raise ConnectionError("rendezvous server closed the connection")
        return self._check(json.loads(line))

    @staticmethod
    def _check(resp):
        if not resp.get("ok"):
            err = resp.get("error", "rendezvous error")
            if err == "timeout":
                raise TimeoutError(err)
            if err == "superseded":
                raise Superseded(err)
            if err == "replaced":
                raise Replaced("a newer incarnation of this rank has joined the rendezvous")
            raise RuntimeError(err)
        return resp.get("value")

    def set(self, key: str, value: bytes) -> None:
        self._call({"op": "set", "key": key, "value": base64.b64encode(value).decode()})

    def get(self, key: str, timeout: Optional[float] = None, scope: str = "", gen: int = -1) -> bytes:
        """Blocks until the key is there. With ``scope`` / ``gen`` (the rendezvous generation the
        caller joined) it raises ``Superseded`` as soon as that generation is over instead of
        waiting out the timeout for a key that will never come."""
        req = {"op": "get", "key": key, "timeout": timeout or self.timeout}
        if scope:
            req.update(scope=scope, gen=gen)
        return base64.b64decode(self._call(req))

    def add(self, key: str, delta: int = 1) -> int:
        return int(self._call({"op": "add", "key": key, "delta": delta}))

    def delete_prefix(self, prefix: str) -> int:
        return int(self._call({"op": "delete_prefix", "prefix": prefix}))

    def join(self, scope: str, rank: int, world: int, timeout: Optional[float] = None,
             ident: Optional[str] = None) -> int:
        """Join the scope's current generation (opening a new one if this rank was already a
        member under another identity, i.e. this process is a restart) and wait until all
        ``world`` ranks are in it. Returns the generation number; keys of this incarnation go
        under ``<scope>/g<gen>/``. Re-joins on its own when a peer's restart supersedes the
        generation while waiting; raises ``Replaced`` when this rank's own successor showed up.
        ``ident`` defaults to pod uid + incarnation + pid of this process."""
        if ident is None:
            ident = (f"{os.environ.get('TFOP_POD_UID', '')}/{os.environ.get('TFOP_INCARNATION', '')}"
                     f"/{os.getpid()}/{self._uuid}")
        end = time.time() + (timeout or self.timeout)
        while True:
            left = max(0.1, end - time.time())
            try:
                return int(self._call({"op": "join", "scope": scope, "rank": rank, "world": world,
                                       "timeout": left, "ident": ident}))
            except Superseded:
                if time.time() >= end:
                    raise TimeoutError("rendezvous join kept being superseded")

    def all_gather(self, prefix: str, rank: int, world: int, value: bytes,
                   timeout: Optional[float] = None, scope: str = "", gen: int = -1) -> List[bytes]:
        """Every rank publishes ``value``; returns all ranks' values in rank order. ``scope`` /
        ``gen``: see ``get``."""
        self.set(f"{prefix}/{rank}", value)
        return [self.get(f"{prefix}/{r}", timeout, scope, gen) for r in range(world)]

    def barrier(self, prefix: str, world: int, timeout: Optional[float] = None, scope: str = "",
                gen: int = -1) -> None:
        n = self.add(f"{prefix}/count", 1)
        round_ = (n - 1) // world       # reusable: round = arrivals // world
        if n % world == 0:
            self.set(f"{prefix}/gen/{round_}", b"1")
        self.get(f"{prefix}/gen/{round_}", timeout, scope, gen)

    # ---- descriptors over the Unix socket --------------------------------------------------
    def _uds(self) -> socket.socket:
        if not self.uds_path:
            raise RuntimeError("no descriptor broker: TFOP_RENDEZVOUS_UDS is not set")
        s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        s.settimeout(self.timeout + 30)
        s.connect(self.uds_path)
        return s

    def fd_put(self, key: str, fd: int) -> None:
        """Hand a file descriptor to the broker (it keeps its own duplicate)."""
        with self._uds() as s:
            socket.send_fds(s, [json.dumps({"op": "fd_put", "key": key}).encode()], [fd])
            self._check(json.loads(s.makefile("rb").readline()))

    def fd_get(self, key: str, timeout: Optional[float] = None, scope: str = "", gen: int = -1) -> int:
        """Receive a duplicate of the descriptor stored under ``key`` (blocks until it is there;
        ``scope`` / ``gen`` as in ``get``). The caller owns the returned descriptor."""
        with self._uds() as s:
            req = {"op": "fd_get", "key": key, "timeout": timeout or self.timeout}
            if scope:
                req.update(scope=scope, gen=gen)
            s.sendall(json.dumps(req).encode())
            msg, fds, _, _ = socket.recv_fds(s, 65536, 1)
            try:
                self._check(json.loads(msg))
            except Exception:
                for fd in fds:
                    os.close(fd)
                raise
            if not fds:
                raise RuntimeError("broker answered without a descriptor")
            return fds[0]

    def close(self) -> None:
        try:
            self._f.close()
            self._sock.close()
        except OSError:
            pass
