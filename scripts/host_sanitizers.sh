#!/bin/bash
# Host libraries under ThreadSanitizer and AddressSanitizer (the reference offers `make DEBUG=1` = ASan only).
# The instrumented library replaces lib/libdllama_b200_host.so for the run and is rebuilt normally afterwards.
#   bash scripts/host_sanitizers.sh          -> prints the number of reports per tool (0 = clean)
set -u
cd "$(dirname "$0")/.."
TESTS="tests/test_json_reader.py tests/test_native_server.py tests/test_native_cpu.py tests/test_text.py tests/test_serving.py tests/test_formats.py"
STD=$(g++ -print-file-name=libstdc++.so)
run() {   # tool, flags, runtime library, options variable
  rm -f /tmp/dllama_san_log*
  CXXFLAGS="-g -O1 $2 -fno-omit-frame-pointer" python -c "import build; build.build_host(force=True)" > /dev/null
  env LD_PRELOAD="$(g++ -print-file-name=$3) $STD" "$4=exitcode=0:log_path=/tmp/dllama_san_log:detect_leaks=0:report_signal_unsafe=0" \
    timeout 1800 python -m pytest $TESTS -q -p no:cacheprovider 2>&1 | tail -n 1
  # reports are blocks separated by "=================="; only those with a frame in this repo's code count
  # (uninstrumented libgomp threads inside libtorch give ThreadSanitizer false positives)
  cat /tmp/dllama_san_log* 2>/dev/null | python3 -c "
import sys
blocks = [b for b in sys.stdin.read().split('==================') if 'SUMMARY' in b]
ours = [b for b in blocks if 'dllama' in b or 'csrc/host' in b]
print('$1: %d report(s) in this repo\'s code, %d elsewhere (torch / OpenMP)' % (len(ours), len(blocks) - len(ours)))
for b in ours[:3]: print(b[:3000])"
}
run ThreadSanitizer -fsanitize=thread libtsan.so TSAN_OPTIONS
run AddressSanitizer -fsanitize=address libasan.so ASAN_OPTIONS
python -c "import build; build.build_host(force=True)" > /dev/null
echo "normal host library restored"
