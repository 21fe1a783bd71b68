#!/usr/bin/env python
"""Generate the checked-in documentation that the reference generates with swagger-codegen and
gen-crd-api-reference-docs:

  docs/sdk/<Client>.md          one page per SDK client (method table + docstrings + signatures)
  docs/sdk/models/<Model>.md    one page per SDK model (attribute / type / JSON key)
  docs/sdk/README.md            index, in the layout of sdk/python/README.md:36-58
  docs/api/kubeflow.org_v1_generated.md   API reference of the four kinds, from the CRD schemas

`hack/update-codegen.sh` runs this; `hack/verify-codegen.sh` runs it into a scratch directory
and diffs.
"""
from __future__ import annotations

import argparse
import inspect
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from tf_operator_b200 import sdk  # noqa: E402
from tf_operator_b200.api import KINDS  # noqa: E402
from tf_operator_b200.api.crd import crd_for  # noqa: E402
from tf_operator_b200.sdk import models as M  # noqa: E402

CLIENTS = ["TFJobClient", "PyTorchJobClient", "MXJobClient", "XGBoostJobClient"]


def _doc(obj) -> str:
    return inspect.cleandoc(obj.__doc__) if obj.__doc__ else ""


def client_page(name: str) -> str:
    cls = getattr(sdk, name)
    methods = [(n, f) for n, f in inspect.getmembers(cls, predicate=inspect.isfunction)
               if not n.startswith("_") or n == "__init__"]
    methods.sort(key=lambda nf: (nf[0] != "__init__", inspect.getsourcelines(nf[1])[1]))
    out = [f"# {name}", "", _doc(cls) or f"Client for `{cls.KIND if hasattr(cls, 'KIND') else name}` objects.", "",
           "| method | description |", "|---|---|"]
    for n, f in methods:
        first = (_doc(f).splitlines() or [""])[0]
        label = name if n == "__init__" else n
        out.append(f"| [{label}](#{label.lower()}) | {first} |")
    out.append("")
    for n, f in methods:
        label = name if n == "__init__" else n
        sig = str(inspect.signature(f)).replace("(self, ", "(").replace("(self)", "()")
        out += [f"## {label}", "", "```python", f"{label}{sig}", "```", "", _doc(f), ""]
    return "\n".join(out).rstrip() + "\n"


def model_names():
    return sorted(n for n in dir(M) if n.startswith("V1") and isinstance(getattr(M, n), type)
                  and issubclass(getattr(M, n), M.Model))


def model_page(name: str) -> str:
    cls = getattr(M, name)
    out = [f"# {name}", "", _doc(cls), "", "## Properties", "",
           "| name | type | JSON key |", "|---|---|---|"]
    for attr, typ in cls.swagger_types.items():
        link = f"[{typ}]({typ}.md)" if typ in model_names() else f"`{typ}`"
        out.append(f"| `{attr}` | {link} | `{cls.attribute_map[attr]}` |")
    out += ["", "`to_dict()`, `to_wire()` (camelCase JSON for the API server), `to_str()`, `==`.", "",
            "[Back to the index](../README.md)", ""]
    return "\n".join(out)


def sdk_index() -> str:
    out = ["# SDK documentation", "",
           "```python", "from tf_operator_b200.sdk import TFJobClient, V1TFJob", "```", "",
           "## Clients", "", "| class | page |", "|---|---|"]
    for cname in CLIENTS:
        out.append(f"| `{cname}` | [{cname}.md]({cname}.md) |")
    out += ["", "## Models", "", "| class | page |", "|---|---|"]
    for m in model_names():
        out.append(f"| `{m}` | [models/{m}.md](models/{m}.md) |")
    out += ["", "A walk-through (create, get, wait, logs, delete) is in "
            "`examples/sdk/tfjob_sdk_walkthrough.py`.", ""]
    return "\n".join(out)


def _schema_rows(schema: dict, prefix: str, rows: list, depth: int = 0) -> None:
    req = set(schema.get("required", []))
    for key, sub in schema.get("properties", {}).items():
        typ = sub.get("type", "object")
        if typ == "array":
            typ = f"[]{sub.get('items', {}).get('type', 'object')}"
        if "enum" in sub:
            typ += " (" + " \\| ".join(str(e) for e in sub["enum"]) + ")"
        desc = " ".join(str(sub.get("description", "")).split())
        rows.append(f"| `{prefix}{key}` | {typ} | {'yes' if key in req else ''} | {desc} |")
        if depth < 3:
            if sub.get("type") == "object" and "properties" in sub:
                _schema_rows(sub, f"{prefix}{key}.", rows, depth + 1)
            elif "additionalProperties" in sub and isinstance(sub["additionalProperties"], dict):
                _schema_rows(sub["additionalProperties"], f"{prefix}{key}.<key>.", rows, depth + 1)


def api_page() -> str:
    out = ["# API reference: kubeflow.org/v1", "",
           "Generated from the CRD schemas (`python -m tf_operator_b200.api.crd`); the pod template "
           "is the full structural core/v1 PodTemplateSpec, as in the reference CRDs (not expanded "
           "here).", ""]
    for kind, mod in KINDS.items():
        schema = crd_for(mod, full=False)["spec"]["versions"][0]["schema"]["openAPIV3Schema"]
        out += [f"## {kind}", "", f"Resource `{mod.INFO.plural}.kubeflow.org`, singular `{mod.INFO.singular}`.", "",
                "| field | type | required | description |", "|---|---|---|---|"]
        rows: list = []
        _schema_rows({"properties": {"spec": schema["properties"]["spec"],
                                     "status": schema["properties"]["status"]}}, "", rows)
        out += rows + [""]
    return "\n".join(out)


def write_all(out_root: Path) -> list:
    written = []

    def put(rel: str, text: str):
        p = out_root / rel
        p.parent.mkdir(parents=True, exist_ok=True)
        p.write_text(text)
        written.append(p)
    for cname in CLIENTS:
        put(f"sdk/{cname}.md", client_page(cname))
    for m in model_names():
        put(f"sdk/models/{m}.md", model_page(m))
    put("sdk/README.md", sdk_index())
    put("api/kubeflow.org_v1_generated.md", api_page())
    return written


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=str(ROOT / "docs"))
    args = ap.parse_args()
    for p in write_all(Path(args.out)):
        print(os.path.relpath(p, ROOT) if str(p).startswith(str(ROOT)) else p)
