"""Launcher-hosted rendezvous store and handle broker.

Replaces the DNS + TCPStore bootstrap of the reference stack (MASTER_ADDR:MASTER_PORT,
pkg/controller.v1/pytorch/pytorch.go:41-65): replicas of a job meet here under
``<namespace>/<job>/...`` keys to publish the CUDA IPC handles of their symmetric buffers
(64-byte blobs) and to barrier. Values are bytes; ``get`` blocks until the key exists.

Wire protocol: one JSON object per line, values base64.
  {"op":"set","key":k,"value":b64} -> {"ok":true}
  {"op":"get","key":k,"timeout":s} -> {"ok":true,"value":b64} | {"ok":false,"error":"timeout"}
  {"op":"add","key":k,"delta":n}   -> {"ok":true,"value":n}
  {"op":"delete_prefix","prefix":p}-> {"ok":true,"value":count}
"""
from __future__ import annotations

import base64
import json
import socket
import socketserver
import threading
import time
from typing import Dict, List, Optional


class RendezvousServer:
    def __init__(self, host: str = "127.0.0.1", port: int = 0):
        self._data: Dict[str, bytes] = {}
        self._counters: Dict[str, int] = {}
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

    @property
    def addr(self) -> str:
        return f"{self.host}:{self.port}"

    def _dispatch(self, req):
        op = req.get("op")
        if op == "set":
            with self._cv:
                self._data[req["key"]] = base64.b64decode(req["value"])
                self._cv.notify_all()
            return {"ok": True}
        if op == "get":
            end = time.time() + float(req.get("timeout", 60))
            with self._cv:
                while req["key"] not in self._data:
                    left = end - time.time()
                    if left <= 0:
                        return {"ok": False, "error": "timeout"}
                    self._cv.wait(left)
                return {"ok": True, "value": base64.b64encode(self._data[req["key"]]).decode()}
        if op == "add":
            with self._cv:
                v = self._counters.get(req["key"], 0) + int(req.get("delta", 1))
                self._counters[req["key"]] = v
                self._data[req["key"]] = str(v).encode()
                self._cv.notify_all()
            return {"ok": True, "value": v}
        if op == "delete_prefix":
            with self._cv:
                ks = [k for k in self._data if k.startswith(req["prefix"])]
                for k in ks:
                    self._data.pop(k, None)
                    self._counters.pop(k, None)
            return {"ok": True, "value": len(ks)}
        return {"ok": False, "error": f"unknown op {op!r}"}

    def start(self) -> "RendezvousServer":
        self._thread = threading.Thread(target=self._srv.serve_forever, daemon=True,
                                        name="tfop-rendezvous")
        self._thread.start()
        return self

    def stop(self) -> None:
        self._srv.shutdown()
        self._srv.server_close()


class RendezvousClient:
    def __init__(self, addr: str, timeout: float = 120.0):
        host, _, port = addr.rpartition(":")
        self._sock = socket.create_connection((host, int(port)), timeout=timeout)
        self._f = self._sock.makefile("rwb")
        self._lock = threading.Lock()
        self.timeout = timeout

    def _call(self, req):
        with self._lock:
            self._f.write(json.dumps(req).encode() + b"\n")
            self._f.flush()
            line = self._f.readline()
        if not line:
            raise ConnectionError("rendezvous server closed the connection")
        resp = json.loads(line)
        if not resp.get("ok"):
            raise TimeoutError(resp.get("error", "rendezvous error")) \
                if resp.get("error") == "timeout" else RuntimeError(resp.get("error"))
        return resp.get("value")

    def set(self, key: str, value: bytes) -> None:
        self._call({"op": "set", "key": key, "value": base64.b64encode(value).decode()})

    def get(self, key: str, timeout: Optional[float] = None) -> bytes:
        return base64.b64decode(self._call({"op": "get", "key": key,
                                            "timeout": timeout or self.timeout}))

    def add(self, key: str, delta: int = 1) -> int:
        return int(self._call({"op": "add", "key": key, "delta": delta}))

    def delete_prefix(self, prefix: str) -> int:
        return int(self._call({"op": "delete_prefix", "prefix": prefix}))

    def all_gather(self, prefix: str, rank: int, world: int, value: bytes,
                   timeout: Optional[float] = None) -> List[bytes]:
        """Every rank publishes ``value``; returns all ranks' values in rank order."""
        self.set(f"{prefix}/{rank}", value)
        return [self.get(f"{prefix}/{r}", timeout) for r in range(world)]

    def barrier(self, prefix: str, world: int, timeout: Optional[float] = None) -> None:
        n = self.add(f"{prefix}/count", 1)
        gen = (n - 1) // world          # reusable: generation = arrivals // world
        if n % world == 0:
            self.set(f"{prefix}/gen/{gen}", b"1")
        self.get(f"{prefix}/gen/{gen}", timeout)

    def close(self) -> None:
        try:
            self._f.close()
            self._sock.close()
        except OSError:
            pass
