#!/usr/bin/env bash
# Same for the legacy TFJob-only binary (reference: scripts/setup-tf-operator.sh).
set -o errexit -o nounset -o pipefail
ROOT=$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)
PORT=${1:-8443}
LOG=${TFOP_LOG:-/tmp/tf-operator.log}
cd "$ROOT"
python -m tf_operator_b200.operator.tf_operator --apiserver-port "$PORT" > "$LOG" 2>&1 &
PID=$!
echo "tf-operator.v1 pid $PID, log $LOG"
for i in $(seq 1 100); do
  if python -c "import urllib.request,sys; urllib.request.urlopen('http://127.0.0.1:$PORT/apis/kubeflow.org/v1/namespaces/default/tfjobs', timeout=2).read()" 2>/dev/null; then
    echo "ready: export TFOP_APISERVER=http://127.0.0.1:$PORT"; exit 0
  fi
  kill -0 "$PID" 2>/dev/null || { echo "operator exited early:" >&2; tail -n 20 "$LOG" >&2; exit 1; }
  sleep 0.2
done
echo "timed out waiting for the operator" >&2; kill "$PID" 2>/dev/null || true; exit 1
