#!/usr/bin/env bash
# compute-sanitizer passes over the single-GPU kernel tests.  Run on a GPU box from the repository root:
#   gpurun --timeout 1500 -- 'bash bench/sanitize.sh > gpurun_out/sanitize.log 2>&1'
# memcheck covers every kernel test; racecheck / synccheck (10-50x slower) cover the shared-memory heavy kernels only.
# The multi-rank P2P kernels are excluded: the tools serialise kernels, and ranks that wait on each other need to run
# concurrently (their single-launch emulation in tests/test_kernels_gpu.py IS covered).
set -u
CS=/usr/local/cuda/bin/compute-sanitizer
PY="python -m pytest -x -q -m gpu"
rc=0
$CS --tool memcheck --error-exitcode 9 --launch-timeout 120 $PY tests/test_kernels_gpu.py tests/test_conv_gpu.py || rc=$?
echo "memcheck rc=$rc"
for tool in racecheck synccheck; do
  $CS --tool $tool --error-exitcode 9 $PY tests/test_kernels_gpu.py -k "bn or maxpool or head or allreduce" || rc=$?
  echo "$tool rc=$rc"
done
exit $rc
