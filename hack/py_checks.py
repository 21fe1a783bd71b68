#!/usr/bin/env python
"""Static checks + CPU unit tests with a junit report (reference: py/kubeflow/tf_operator/
py_checks.py - pylint over every .py file, then the py unit tests, each as a junit case).

    python hack/py_checks.py --junit_path /tmp/junit_pychecks.xml [--no-tests]

Checks: every tracked .py byte-compiles; no file references an undefined top-level name that
`compile` can catch; then `pytest -m "not gpu"`.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tf_operator_b200.testing import util  # noqa: E402

SKIP_DIRS = {".git", "gpurun_out", "baseline", "__pycache__", "build", "dist"}


def python_files():
    for base, dirs, files in os.walk(ROOT):
        dirs[:] = [d for d in dirs if d not in SKIP_DIRS]
        for f in files:
            if f.endswith(".py"):
                yield os.path.join(base, f)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--junit_path", default="")
    ap.add_argument("--no-tests", action="store_true")
    args = ap.parse_args()
    cases = []
    for path in sorted(python_files()):
        tc = util.TestCase("py_compile", os.path.relpath(path, ROOT))
        start = time.time()
        try:
            with open(path, "rb") as f:
                compile(f.read(), path, "exec")       # syntax check without writing .pyc files
        except SyntaxError as e:
            tc.failure = f"{type(e).__name__}: {e}"
        tc.time = time.time() - start
        cases.append(tc)
    if not args.no_tests:
        tc = util.TestCase("pytest", "not_gpu")
        start = time.time()
        res = subprocess.run([sys.executable, "-m", "pytest", "tests", "-x", "-q", "-m", "not gpu"],
                             cwd=ROOT, capture_output=True, text=True)
        tc.time = time.time() - start
        if res.returncode != 0:
            tc.failure = res.stdout[-4000:]
        cases.append(tc)
    failed = [c for c in cases if c.failure]
    if args.junit_path:
        util.create_junit_xml_file(cases, args.junit_path)
    print(f"{len(cases)} checks, {len(failed)} failed")
    for c in failed:
        print("FAIL", c.class_name, c.name, "\n", c.failure[:800])
    return 1 if failed else 0


if __name__ == "__main__":
    sys.exit(main())
