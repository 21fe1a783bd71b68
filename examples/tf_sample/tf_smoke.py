#!/usr/bin/env python
"""Smoke test of the cluster wiring — the reference's tf_smoke (examples/tf_sample/tf_sample/
tf_smoke.py:36-142): every non-chief task just serves; the chief (master/chief, or worker 0 when
there is neither) asks EVERY task of the cluster, including itself, to multiply two 10x10
integer matrices and checks the answers, retrying while peers are still starting."""
import json
import os
import socket
import socketserver
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402


class Handler(socketserver.StreamRequestHandler):
    def handle(self):
        for line in self.rfile:
            req = json.loads(line)
            if req.get("op") == "matmul":
                a, b = req["a"], req["b"]
                n = len(a)
                c = [[sum(a[i][k] * b[k][j] for k in range(n)) for j in range(n)] for i in range(n)]
                self.wfile.write(json.dumps({"c": c}).encode() + b"\n")
            elif req.get("op") == "shutdown":
                self.wfile.write(b"{}\n")
                threading.Thread(target=self.server.shutdown, daemon=True).start()
                return


def serve(endpoint):
    host, _, port = endpoint.rpartition(":")

    class S(socketserver.ThreadingTCPServer):
        allow_reuse_address = True
        daemon_threads = True
    srv = S((host, int(port)), Handler)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    return srv


def call(endpoint, req, retries=100):
    host, _, port = endpoint.rpartition(":")
    last = None
    for _ in range(retries):          # @retrying.retry around grpc Unavailable in the reference
        try:
            with socket.create_connection((host, int(port)), timeout=5) as s:
                f = s.makefile("rwb")
                f.write(json.dumps(req).encode() + b"\n")
                f.flush()
                return json.loads(f.readline())
        except OSError as e:
            last = e
            time.sleep(0.1)
    raise ConnectionError(f"{endpoint}: {last}")


def main():
    ctx = ReplicaContext.from_env()
    if ctx.kind != "TFJob":
        print("no TF_CONFIG: running locally")
        return 0
    srv = serve(ctx.my_endpoint())
    print(f"{ctx.task_type}:{ctx.task_index} serving on {ctx.my_endpoint()}", flush=True)
    if not ctx.is_chief:
        threading.Event().wait()                      # server.join()
        return 0
    a = [[(i + j) % 7 for j in range(10)] for i in range(10)]
    b = [[(i * j) % 5 for j in range(10)] for i in range(10)]
    want = [[sum(a[i][k] * b[k][j] for k in range(10)) for j in range(10)] for i in range(10)]
    for role in sorted(ctx.cluster):
        for idx, ep in enumerate(ctx.resolved(role)):
            got = call(ep, {"op": "matmul", "a": a, "b": b})["c"]
            assert got == want, (role, idx)
            print(f"/job:{role}/task:{idx} ok", flush=True)
    print("all tasks answered", flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
