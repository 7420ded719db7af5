#!/bin/bash
# The memcheck stand-in that needs no GPU: the whole library compiled as host C++ under AddressSanitizer
# (tests/emu_lib/build.py --asan; "device" memory is heap memory) and a selection of the device tests plus the sharded
# solves with ranks as host threads run on it.  Any out-of-bounds access of a kernel, a copy or the host driver stops the
# run with a report that names the source line.     bash scripts/emulated_asan.sh > profiles/r02_emulated_asan.log 2>&1
set -u
cd "$(dirname "$0")/.."
python tests/emu_lib/build.py --asan || exit 1
export LD_PRELOAD=$(gcc -print-file-name=libasan.so)
export ASAN_OPTIONS=detect_leaks=0:halt_on_error=1
export PDLP_B200_LIBRARY=$PWD/tests/_build/libpdlp_b200_emu_asan.so
export PDLP_B200_NCCL_LIBRARY=$PWD/tests/_build/libfake_nccl.so
python -m pytest -q -m gpu -p no:cacheprovider --timeout 900 tests/test_zy_late_additions.py tests/test_python_surface.py tests/test_presolve.py \
  tests/test_reference_units.py tests/test_gpu_trial_steps.py tests/test_gpu_forced_termination.py tests/test_reference_pdhg_behaviour.py \
  -k "not concurrent and not plain_c and not reports_on_the_original" 2>&1 | tail -5
for cfg in "2 peer random" "8 peer flow" "2 nccl random"; do
  python tests/emu_lib/sharded_threads.py $cfg 2>&1 | tail -1 | cut -c1-400
done
echo "done (a report would have stopped the run: halt_on_error=1)"
