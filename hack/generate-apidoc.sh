#!/usr/bin/env bash
# API reference only (reference: hack/generate-apidoc.sh drives gen-crd-api-reference-docs).
set -o errexit -o nounset -o pipefail
ROOT=$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)
cd "$ROOT"
python - <<'PY'
import sys
sys.path.insert(0, "hack")
from pathlib import Path
import gen_docs
p = Path("docs/api/kubeflow.org_v1_generated.md")
p.parent.mkdir(parents=True, exist_ok=True)
p.write_text(gen_docs.api_page())
print(p)
PY
