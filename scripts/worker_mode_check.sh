#!/bin/bash
# Root + worker processes over the TCP bootstrap (`dllama worker --port` / `--workers host:port`), one GPU
# each, on one box: the reference's deployment shape.  Generates a tiny TP-capable model + tokenizer,
# starts W workers, runs `dllama inference` on the root and checks that it printed the benchmark block.
# Usage: W=1 bash scripts/worker_mode_check.sh        (needs W+1 GPUs)
set -u
cd "$(dirname "$0")/.."
W="${W:-1}"
T=$(mktemp -d)
python - "$T" <<'PY'
import sys
from dllama_b200.formats import mfile, tfile
d = sys.argv[1]
mfile.write_random_model(d + "/tiny.m", "tiny-tp8", seed=3)
tfile.write_synthetic_tokenizer(d + "/tiny.t", 1024)
PY
PIDS=()
ARGS=()
for (( w = 0; w < W; w++ )); do
  port=$((9999 - w))
  ./dllama worker --port "$port" --gpu-index $((w + 1)) > "$T/worker_$w.log" 2>&1 &
  PIDS+=($!)
  ARGS+=("127.0.0.1:$port")
done
sleep 3
timeout 240 ./dllama inference --model "$T/tiny.m" --tokenizer "$T/tiny.t" --buffer-float-type q80 --prompt "hello world the of and" \
  --steps 24 --max-seq-len 256 --gpu-index 0 --seed 7 --workers "${ARGS[@]}" > "$T/root.log" 2>&1
rc=$?
for p in "${PIDS[@]}"; do kill "$p" 2>/dev/null; done
tail -n 12 "$T/root.log"
for (( w = 0; w < W; w++ )); do echo "--- worker $w"; tail -n 3 "$T/worker_$w.log"; done
if [ $rc -eq 0 ] && grep -q "Prediction" "$T/root.log" && grep -q "Socket\[$((W - 1))\]" "$T/root.log"; then echo "WORKER_MODE PASS world=$((W + 1))"; else echo "WORKER_MODE FAIL rc=$rc"; exit 1; fi
