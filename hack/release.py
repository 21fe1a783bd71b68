#!/usr/bin/env python
"""Cut a release artefact (reference: py/kubeflow/tf_operator/release.py builds and pushes the
operator image and a helm chart from a git commit; on one box the artefact is a source tarball
with the built sm_100a library, the generated manifests and a checksum).

    python hack/release.py --out dist [--no-build]
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import subprocess
import sys
import tarfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tf_operator_b200 import version  # noqa: E402

INCLUDE = ["tf_operator_b200", "include", "manifests", "examples", "docs", "hack", "scripts/setup-training-operator.sh",
           "scripts/setup-tf-operator.sh", "bench.py", "__graft_entry__.py", "README.md", "DESIGN.md",
           "INTEGRATION.md", "pytest.ini", "tests"]
EXCLUDE_PARTS = {"__pycache__", "obj"}


def git_rev() -> str:
    try:
        return subprocess.run(["git", "rev-parse", "--short", "HEAD"], cwd=ROOT, capture_output=True,
                              text=True, check=True).stdout.strip()
    except Exception:  # noqa: BLE001
        return "unknown"


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "dist"))
    ap.add_argument("--no-build", action="store_true", help="do not (re)build libtfop_b200.so first")
    args = ap.parse_args()
    if not args.no_build:
        from tf_operator_b200.ops import _build
        _build.build()
    subprocess.run(["bash", os.path.join(ROOT, "hack", "verify-codegen.sh")], check=True)
    os.makedirs(args.out, exist_ok=True)
    ver = getattr(version, "VERSION", "0.0.0")
    name = f"tf_operator_b200-{ver}+{git_rev()}"
    tar_path = os.path.join(args.out, name + ".tar.gz")

    def keep(ti: tarfile.TarInfo):
        parts = set(ti.name.split("/"))
        return None if parts & EXCLUDE_PARTS else ti
    with tarfile.open(tar_path, "w:gz") as tar:
        for rel in INCLUDE:
            p = os.path.join(ROOT, rel)
            if os.path.exists(p):
                tar.add(p, arcname=f"{name}/{rel}", filter=keep)
    sha = hashlib.sha256(open(tar_path, "rb").read()).hexdigest()
    meta = {"name": name, "version": ver, "git": git_rev(), "sha256": sha,
            "built_at": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()),
            "arch": "sm_100a", "artifact": os.path.basename(tar_path)}
    with open(os.path.join(args.out, name + ".json"), "w") as f:
        json.dump(meta, f, indent=1)
    print(json.dumps(meta))
    return 0


if __name__ == "__main__":
    sys.exit(main())
