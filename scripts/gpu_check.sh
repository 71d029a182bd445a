#!/bin/bash
# Run every GPU test function in its own process (a faulting kernel must not poison the CUDA context
# of the other tests), then a short bench.  Output -> gpurun_out/.
# SANITIZE=memcheck|racecheck|synccheck|initcheck runs every test under compute-sanitizer with that tool.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OUT=gpurun_out/gpu_check.log
: > $OUT
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv >> $OUT 2>&1
FILES="${@:-tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_f32_engine.py tests/test_zz_no_alloc.py}"
for f in $FILES; do
  for t in $(python -m pytest $f --collect-only -q -m gpu 2>/dev/null | grep "::" | sed 's/\[.*//' | sort -u); do
    echo "=== $t" >> $OUT
    if [ -n "${SANITIZE:-}" ]; then
      timeout 1200 compute-sanitizer --tool "$SANITIZE" --error-exitcode 1 python -m pytest "$t" -q -m gpu -x --no-header -p no:cacheprovider 2>&1 | tail -n 40 >> $OUT
    else
      timeout 300 python -m pytest "$t" -q -m gpu -x --no-header -p no:cacheprovider 2>&1 | tail -n 25 >> $OUT
    fi
    echo "=== exit $? $t" >> $OUT
  done
done
grep -E "^=== exit|passed|failed|error" $OUT | tail -n 60
