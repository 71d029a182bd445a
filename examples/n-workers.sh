#!/bin/bash
# Start / stop N worker processes (one per extra GPU of this box) for a root started with
#   ./dllama inference ... --workers 127.0.0.1:9999 127.0.0.1:9998 ...
# Usage:  W=7 bash examples/n-workers.sh start      W=7 bash examples/n-workers.sh stop
# Worker w listens on port 9999-w and uses GPU w+1 (GPU 0 is the root's).
# Equivalent without workers:  torchrun --nproc-per-node N dllama ...
# Without GPUs:  DEVICE=cpu NTHREADS=4 W=1 bash examples/n-workers.sh start   (workers of the host backend; they get
# their weight slices from the root over TCP, so they need no model file)
set -u
cd "$(dirname "$0")/.."
W="${W:-3}"
PIDS=examples/.workers.pids
case "${1:-}" in
  start)
    : > "$PIDS"
    for (( w = 0; w < W; w++ )); do
      port=$((9999 - w))
      mkdir -p "examples/dllama_worker_$w"
      if [ "${DEVICE:-auto}" = "cpu" ]; then
        nohup ./dllama worker --port "$port" --device cpu --nthreads "${NTHREADS:-0}" > "examples/dllama_worker_$w/log.txt" 2>&1 &
      else
        nohup ./dllama worker --port "$port" --gpu-index $((w + 1)) > "examples/dllama_worker_$w/log.txt" 2>&1 &
      fi
      echo $! >> "$PIDS"
      echo "Started worker $w on port $port (GPU $((w + 1)), pid $!)"
    done
    sleep 2
    ;;
  stop)
    if [ -f "$PIDS" ]; then
      while read -r pid; do kill "$pid" 2>/dev/null && echo "Stopped $pid"; done < "$PIDS"
      rm -f "$PIDS"
    fi
    ;;
  *) echo "Usage: W=<n> $0 [start|stop]"; exit 1 ;;
esac
