#!/usr/bin/env python
"""Fake workload for deterministic control-plane tests: an HTTP server on the replica's
address that echoes its environment and exits with a chosen code on demand.

Same three endpoints as the reference's test/test-server/test_app.py:19-82:
  /tfconfig            -> the TF_CONFIG this replica received
  /runconfig           -> the tf.estimator.RunConfig fields TensorFlow would derive from it
  /exit?exitCode=N     -> the process exits with N (fault injection)
Served with the standard library (Flask is not installed here).
"""
import json
import os
import sys
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from urllib.parse import parse_qs, urlparse

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from tf_operator_b200.runtime.runconfig import run_config_from_tf_config  # noqa: E402


class Handler(BaseHTTPRequestHandler):
    def log_message(self, *a):
        pass

    def _send(self, body: str, ctype="text/plain"):
        data = body.encode()
        self.send_response(200)
        self.send_header("Content-Type", ctype)
        self.send_header("Content-Length", str(len(data)))
        self.end_headers()
        self.wfile.write(data)

    def do_GET(self):  # noqa: N802
        u = urlparse(self.path)
        if u.path == "/":
            return self._send("hello world")
        if u.path == "/tfconfig":
            return self._send(os.environ.get("TF_CONFIG", ""))
        if u.path == "/runconfig":
            cfg = run_config_from_tf_config(os.environ.get("TF_CONFIG", ""))
            return self._send(json.dumps(cfg), "application/json")
        if u.path == "/env":
            return self._send(json.dumps(dict(os.environ)), "application/json")
        if u.path == "/exit":
            code = int(parse_qs(u.query).get("exitCode", ["0"])[0])
            self._send(f"exit {code}")
            threading.Thread(target=lambda: os._exit(code), daemon=True).start()
            return None
        self.send_response(404)
        self.end_headers()
        return None


def main():
    host = os.environ.get("TFOP_POD_IP", "127.0.0.1")
    port = int(os.environ.get("PORT", "2222"))
    srv = ThreadingHTTPServer((host, port), Handler)
    print(f"test-server listening on {host}:{port}", flush=True)
    srv.serve_forever()


if __name__ == "__main__":
    main()
