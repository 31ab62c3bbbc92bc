#!/usr/bin/env bash
# compute-sanitizer over the single-GPU kernel tests, ONE tool per invocation (and per gpurun call):
#   gpurun --timeout 1200 -- 'bash bench/sanitize.sh memcheck  > gpurun_out/sanitize_memcheck.log 2>&1'
#   gpurun --timeout 1200 -- 'bash bench/sanitize.sh racecheck > gpurun_out/sanitize_racecheck.log 2>&1'
# memcheck covers every kernel test; racecheck / synccheck / initcheck (10-50x slower) cover the shared-memory heavy
# kernels only.  The multi-rank P2P kernels are excluded: the tools serialise kernels, and ranks that wait on each other
# need to run concurrently (their single-launch emulation in tests/test_kernels_gpu.py IS covered).
# If a bounded mbarrier wait trips under the tool's slow-down, rebuild first with
#   DMPI_NVCC_EXTRA="-DDMPI_MBAR_TIMEOUT_NS=0" python build.py --force      (csrc/common/ptx.cuh)
set -u
tool="${1:-memcheck}"
CS=/usr/local/cuda/bin/compute-sanitizer
PY="python -m pytest -x -q -m gpu -p no:cacheprovider"
export DMPI_COMM_TIMEOUT_S=0
rc=0
if [ "$tool" = "memcheck" ]; then
  # SANITIZE_SELECT narrows the run (pytest -k expression) when GPU minutes are short
  $CS --tool memcheck --error-exitcode 9 --launch-timeout 120 $PY tests/test_kernels_gpu.py tests/test_conv_gpu.py \
      ${SANITIZE_SELECT:+-k "$SANITIZE_SELECT"} || rc=$?
else
  $CS --tool "$tool" --error-exitcode 9 $PY tests/test_kernels_gpu.py -k "bn or maxpool or head or allreduce" || rc=$?
fi
echo "$tool rc=$rc"
exit $rc
