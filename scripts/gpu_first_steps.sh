#!/bin/bash
# What to run first on a B200 box after host-side changes that were made without a GPU (one gpurun call each;
# everything lands in gpurun_out/, copy what should be judged into profiles/).
#
#   gpurun --timeout 1500 -- 'bash scripts/gpu_first_steps.sh tests'      GPU suite, stops at the first failure
#   gpurun --timeout 900  -- 'bash scripts/gpu_first_steps.sh bench'      smoke() + bench.py N=1 (own arm)
#   gpurun --timeout 1200 -- 'bash scripts/gpu_first_steps.sh serving'    dllama-api over HTTP: native vs Python front end,
#                                                                         greedy and sampled, 32 streaming users
#   gpurun --timeout 900  -- 'bash scripts/gpu_first_steps.sh longprompt' a ~4000-token prompt (1300 words) alone through dllama-api
#   gpurun --gpus 8 --timeout 2400 -- 'bash scripts/gpu_first_steps.sh 405b'   Llama-3.1-405B Q40 at TP=8 (238 GB file)
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
case "${1:-tests}" in
  tests)
    timeout 1400 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 40 | tee gpurun_out/gpu_tests.log
    ;;
  bench)
    timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 3 | tee gpurun_out/smoke.log
    timeout 600 python bench.py --users 32 2> gpurun_out/bench_n1.err | tee gpurun_out/bench_n1.json
    ;;
  serving)
    for backend in native python; do
      for temp in 0 0.8; do
        timeout 280 python scripts/load_test_api.py --users 32 --max-tokens 64 --backend $backend --temperature $temp \
          2>&1 | tail -n 2 | tee -a gpurun_out/load_test_api_${backend}.log
      done
    done
    ;;
  longprompt)
    timeout 600 python scripts/load_test_api.py --users 1 --max-tokens 16 --prompt-words 1300 --max-batch 256 --max-seq-len 8192 --rounds 2 \
      2>&1 | tail -n 2 | tee gpurun_out/load_test_api_longprompt.log
    ;;
  405b)
    export DLLAMA_FAST_WRITER=1
    N=$(nvidia-smi -L | wc -l)
    timeout 2300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29531 \
      bench.py --gpus "$N" --model llama3.1-405b --steps 16 --warmup 3 --users 32 --prompt-len 2048 \
      2> gpurun_out/bench_405b.err | tee gpurun_out/bench_405b_tp${N}.json
    tail -n 5 gpurun_out/bench_405b.err
    ;;
  *) echo "usage: $0 {tests|bench|serving|longprompt|405b}"; exit 1 ;;
esac
