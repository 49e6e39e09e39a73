#!/usr/bin/env bash
# Bring up one job on this machine and collect its metrics (role of the reference's terraform/deploy.sh:22-65 +
# ECS services in terraform/main.tf:384-435: "start 1 server and W workers, wait, read METRICS_JSON").
#
#   deploy/run_box.sh gpu  <W> <sync|async> [extra launch args]     # W GPUs of this box, NVLink data path
#   deploy/run_box.sh grpc <W> <sync|async> [extra worker args]     # 1 gRPC server + W CPU workers on localhost
#
# Results land in out/<mode>_<W>workers/ in the reference's result-file schema.
set -euo pipefail
kind=${1:?gpu|grpc}; W=${2:?workers}; mode=${3:?sync|async}; shift 3
here=$(cd "$(dirname "$0")/.." && pwd); cd "$here"
out="out/${mode}_${W}workers"; mkdir -p "$out"
if [ "$kind" = gpu ]; then
  exec python -m dps_b200.launch --gpus "$W" -- --mode "$mode" --workers "$W" --metrics-dir "$out" "$@"
fi
port=${SERVER_PORT:-8000}
SERVER_MODE=$mode TOTAL_WORKERS_EXPECTED=$W SERVER_PORT=$port \
  python -m dps_b200.rpc.servicer --metrics-dir "$out" > "$out/server.log" 2>&1 &
server=$!
trap 'kill $server 2>/dev/null || true' EXIT
pids=()
for i in $(seq 1 "$W"); do
  PARAMETER_SERVER_ADDRESS=localhost:$port python -m dps_b200.rpc.client --worker-name "worker-$i" \
    --metrics-dir "$out" "$@" > "$out/worker-$i.log" 2>&1 &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
sleep 1                                   # the server prints its final block when the last worker leaves
kill $server 2>/dev/null || true
python -m dps_b200.utils.charts collect --metrics-dir "$out" --name "${mode}_${W}workers" --output "$out/result.json"
