#!/usr/bin/env bash
# Regenerate everything that is derived from the API types (reference: hack/update-codegen.sh
# :76-80 runs k8s code-generator for the Go client, deepcopy and OpenAPI; here the derived
# artefacts are the CRD manifests and the SDK / API documentation).
set -o errexit -o nounset -o pipefail
ROOT=$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)
cd "$ROOT"
python -m tf_operator_b200.api.crd manifests/base
python hack/gen_docs.py --out docs
echo "codegen: CRDs -> manifests/base, docs -> docs/{sdk,api}"
