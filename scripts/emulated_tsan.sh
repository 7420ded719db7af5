#!/bin/bash
# The racecheck stand-in that needs no GPU: the whole library as host C++ in THREAD mode (one host thread per
# CUDA thread) under ThreadSanitizer, and six whole solves on it (tests/emu_lib/tsan_solves.cc).  A data race between the
# threads of a CTA, or in the host driver, is printed as "WARNING: ThreadSanitizer".  Minutes per solve: an emulated
# launch creates 256 threads.       bash scripts/emulated_tsan.sh > profiles/r02_emulated_tsan.log 2>&1
set -u
cd "$(dirname "$0")/.."
python tests/emu_lib/build.py --tsan || exit 1
mkdir -p tests/_build/emu_callers
g++ -std=c++17 -g -O1 -fsanitize=thread -Iinclude tests/emu_lib/tsan_solves.cc -Ltests/_build -lpdlp_b200_emu_tsan -Wl,-rpath,$PWD/tests/_build \
  -o tests/_build/emu_callers/tsan_solves || exit 1
TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" ./tests/_build/emu_callers/tsan_solves 2>&1 | tee tests/_build/emu_callers/tsan_solves.out | grep -v "^\s*#" | head -200
echo "exit ${PIPESTATUS[0]}; ThreadSanitizer warnings: $(grep -c 'WARNING: ThreadSanitizer' tests/_build/emu_callers/tsan_solves.out)"
