#!/usr/bin/env bash
# Fail when the checked-in generated files differ from what the API types produce now
# (reference: hack/verify-codegen.sh - regenerate into a scratch tree and diff).
set -o errexit -o nounset -o pipefail
ROOT=$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)
TMP=$(mktemp -d)
trap 'rm -rf "$TMP"' EXIT
cd "$ROOT"
python -m tf_operator_b200.api.crd "$TMP/manifests" > /dev/null
python hack/gen_docs.py --out "$TMP/docs" > /dev/null
rc=0
for f in "$TMP"/manifests/*.yaml; do
  diff -u "manifests/base/$(basename "$f")" "$f" || rc=1
done
diff -ru docs/sdk "$TMP/docs/sdk" || rc=1
diff -ru docs/api "$TMP/docs/api" || rc=1
if [ $rc -ne 0 ]; then
  echo "generated files are out of date: run hack/update-codegen.sh" >&2
  exit 1
fi
echo "generated files are up to date"
