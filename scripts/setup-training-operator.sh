#!/usr/bin/env bash
# Bring up the all-in-one training operator and wait until it is ready (reference:
# scripts/setup-training-operator.sh = kustomize edit set image + kubectl apply + wait for the
# deployment and the four CRDs). On one box "deploy" = start the manager process with its
# in-memory API server + REST facade + launcher, then poll /readyz and the CRD endpoints.
#
#   scripts/setup-training-operator.sh [PORT]     # default 8443; prints the PID and the URL
#   export TFOP_APISERVER=http://127.0.0.1:8443   # what the SDK and examples/run_job.py read
set -o errexit -o nounset -o pipefail
ROOT=$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)
PORT=${1:-8443}
LOG=${TFOP_LOG:-/tmp/training-operator.log}
cd "$ROOT"
python -m tf_operator_b200.operator.training_operator --apiserver-port "$PORT" \
  --enable-scheme tfjob --enable-scheme pytorchjob --enable-scheme mxjob --enable-scheme xgboostjob \
  > "$LOG" 2>&1 &
PID=$!
echo "training-operator pid $PID, log $LOG"
for i in $(seq 1 100); do
  if python - "$PORT" <<'PY' 2>/dev/null
import sys, urllib.request
base = f"http://127.0.0.1:{sys.argv[1]}"
for plural in ("tfjobs", "pytorchjobs", "mxjobs", "xgboostjobs"):
    urllib.request.urlopen(f"{base}/apis/apiextensions.k8s.io/v1/customresourcedefinitions/{plural}.kubeflow.org", timeout=2).read()
PY
  then
    echo "ready: export TFOP_APISERVER=http://127.0.0.1:$PORT"
    exit 0
  fi
  if ! kill -0 "$PID" 2>/dev/null; then
    echo "operator exited early:" >&2; tail -n 20 "$LOG" >&2; exit 1
  fi
  sleep 0.2
done
echo "timed out waiting for the operator" >&2
kill "$PID" 2>/dev/null || true
exit 1
