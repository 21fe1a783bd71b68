"""Parameter-server transport for the TFJob PS/Worker pattern, CPU path.

The reference's dist-mnist job (examples/v1/dist-mnist/dist_mnist.py:161-208) places variables
on /job:ps and has every worker pull them and push gradients over TF gRPC, with the optimizer
applied on the PS, asynchronously. This is that protocol over plain TCP for the CPU plumbing
config (BASELINE.json config 1): variables are round-robined over the PS shards, ``pull``
returns the shard's current values, ``push`` applies Adam on the shard and bumps the shared
global step. The GPU path of the same pattern (P2P embedding gather + fused Adagrad on the
owning shard) is ops/ps_embedding.py + csrc/ps.cu.
"""
from __future__ import annotations

import io
import pickle
import socket
import socketserver
import struct
import threading
from typing import Dict, List, Optional, Tuple

import torch


def _send(sock: socket.socket, obj) -> None:
    buf = io.BytesIO()
    torch.save(obj, buf)
    data = buf.getvalue()
    sock.sendall(struct.pack("!Q", len(data)) + data)


def _recv(sock: socket.socket):
    hdr = b""
    while len(hdr) < 8:
        chunk = sock.recv(8 - len(hdr))
        if not chunk:
            raise ConnectionError("peer closed")
        hdr += chunk
    (n,) = struct.unpack("!Q", hdr)
    data = bytearray()
    while len(data) < n:
        chunk = sock.recv(min(1 << 20, n - len(data)))
        if not chunk:
            raise ConnectionError("peer closed")
        data += chunk
    # weights_only: the port is reachable by any local process; the payloads are plain
    # containers of tensors, numbers and strings, so the restricted unpickler is enough
    return torch.load(io.BytesIO(bytes(data)), weights_only=True)


class ParameterServer:
    """One PS shard: owns a dict of variables and their Adam state."""

    def __init__(self, host: str, port: int, lr: float = 0.01, betas=(0.9, 0.999), eps: float = 1e-8):
        self.vars: Dict[str, torch.Tensor] = {}
        self.m: Dict[str, torch.Tensor] = {}
        self.v: Dict[str, torch.Tensor] = {}
        self.t: Dict[str, int] = {}
        self.lr, self.betas, self.eps = lr, betas, eps
        self.global_step = 0
        self.pushes = 0
        self._lock = threading.Lock()
        self._done = threading.Event()
        outer = self

        class Handler(socketserver.BaseRequestHandler):
            def handle(self):
                try:
                    while True:
                        req = _recv(self.request)
                        _send(self.request, outer._dispatch(req))
                except (ConnectionError, OSError):
                    return

        class Server(socketserver.ThreadingTCPServer):
            allow_reuse_address = True
            daemon_threads = True

        self._srv = Server((host, port), Handler)

    def _dispatch(self, req):
        op = req["op"]
        with self._lock:
            if op == "init":          # first writer wins (chief initialises the variables)
                for k, t in req["vars"].items():
                    if k not in self.vars:
                        self.vars[k] = t.clone()
                        self.m[k] = torch.zeros_like(t)
                        self.v[k] = torch.zeros_like(t)
                        self.t[k] = 0
                return {"ok": True}
            if op == "pull":
                return {"vars": {k: self.vars[k].clone() for k in req["names"] if k in self.vars},
                        "global_step": self.global_step}
            if op == "push":
                b1, b2 = self.betas
                for k, g in req["grads"].items():
                    self.t[k] += 1
                    self.m[k].mul_(b1).add_(g, alpha=1 - b1)
                    self.v[k].mul_(b2).addcmul_(g, g, value=1 - b2)
                    mh = self.m[k] / (1 - b1 ** self.t[k])
                    vh = self.v[k] / (1 - b2 ** self.t[k])
                    self.vars[k].addcdiv_(mh, vh.sqrt().add_(self.eps), value=-self.lr)
                self.pushes += 1
                if req.get("count_step"):
                    self.global_step += 1
                return {"global_step": self.global_step}
            if op == "step":
                return {"global_step": self.global_step}
            if op == "shutdown":
                self._done.set()
                return {"ok": True}
        return {"error": f"unknown op {op}"}

    def serve_forever(self):
        """``server.join()``: PS replicas never exit on their own — which is why
        cleanPodPolicy=Running exists."""
        t = threading.Thread(target=self._srv.serve_forever, daemon=True)
        t.start()
        self._done.wait()
        self._srv.shutdown()

    def start(self):
        t = threading.Thread(target=self._srv.serve_forever, daemon=True)
        t.start()
        return self

    def stop(self):
        self._done.set()
        self._srv.shutdown()
        self._srv.server_close()


class PSClient:
    """Worker-side handle on all PS shards; variables are round-robined by sorted name."""

    def __init__(self, endpoints: List[str], timeout: float = 60.0, retries: int = 100):
        import time
        self.socks: List[socket.socket] = []
        for ep in endpoints:
            host, _, port = ep.rpartition(":")
            last = None
            for _ in range(retries):
                try:
                    s = socket.create_connection((host, int(port)), timeout=timeout)
                    s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                    self.socks.append(s)
                    break
                except OSError as e:      # PS not up yet: retry like tf_smoke's @retrying
                    last = e
                    time.sleep(0.1)
            else:
                raise ConnectionError(f"cannot reach parameter server {ep}: {last}")
        self.placement: Dict[str, int] = {}

    def place(self, names: List[str]) -> None:
        for i, n in enumerate(sorted(names)):
            self.placement[n] = i % len(self.socks)

    def _call(self, shard: int, req):
        _send(self.socks[shard], req)
        return _recv(self.socks[shard])

    def init(self, variables: Dict[str, torch.Tensor]) -> None:
        self.place(list(variables))
        for shard in range(len(self.socks)):
            mine = {k: v for k, v in variables.items() if self.placement[k] == shard}
            self._call(shard, {"op": "init", "vars": mine})

    def pull(self, names: List[str]) -> Tuple[Dict[str, torch.Tensor], int]:
        if not self.placement:
            self.place(names)
        out: Dict[str, torch.Tensor] = {}
        step = 0
        for shard in range(len(self.socks)):
            mine = [n for n in names if self.placement[n] == shard]
            if not mine and shard != 0:
                continue
            resp = self._call(shard, {"op": "pull", "names": mine})
            out.update(resp["vars"])
            if shard == 0:
                step = resp["global_step"]
        return out, step

    def push(self, grads: Dict[str, torch.Tensor]) -> int:
        step = 0
        for shard in range(len(self.socks)):
            mine = {k: g for k, g in grads.items() if self.placement[k] == shard}
            resp = self._call(shard, {"op": "push", "grads": mine, "count_step": shard == 0})
            if shard == 0:
                step = resp["global_step"]
        return step

    def global_step(self) -> int:
        return self._call(0, {"op": "step"})["global_step"]

    def close(self):
        for s in self.socks:
            try:
                s.close()
            except OSError:
                pass
