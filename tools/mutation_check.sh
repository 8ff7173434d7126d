#!/bin/bash
# Does tests/test_sky_gates_random_gpu.py have teeth?  Builds three deliberately wrong librc variants (a column region one
# step too deep; sky gates without their slack; drop gates that count the drop to the near edge of a ring only) and runs the randomised test against each on the GPU box: both must FAIL.
#   here:        tools/mutation_check.sh build
#   on the box:  tools/mutation_check.sh run   (under gpurun; writes gpurun_out/mutation_check.log)
# Without a GPU: KHOST_DEFINES=-DRC_MUTATE_DROP python -m pytest tests/test_kernel_source_cpu.py -x -q  (the same three mutants in
# the host build of kernels.cu: each fails at the first frame).
set -u
cd "$(dirname "$0")/.."
if [ "${1:-}" = build ]; then
  tools/build_variants.sh mut_column "-DRC_MUTATE_COLUMN" mut_gate "-DRC_MUTATE_GATE" mut_drop "-DRC_MUTATE_DROP" > /dev/null 2>&1
  ls -la raytracing_test_b200/variants/librc_mut_*.so
  exit 0
fi
mkdir -p gpurun_out
: > gpurun_out/mutation_check.log
for m in mut_column mut_gate mut_drop; do
  RC_LIBRC=$PWD/raytracing_test_b200/variants/librc_$m.so python -m pytest tests/test_sky_gates_random_gpu.py -m gpu -x -q > gpurun_out/mutation_$m.log 2>&1
  rc=$?
  echo "$m: pytest exit code $rc ($([ $rc -ne 0 ] && echo 'caught' || echo 'NOT CAUGHT'))" >> gpurun_out/mutation_check.log
  grep -E "AssertionError|seed [0-9]+" gpurun_out/mutation_$m.log | head -3 >> gpurun_out/mutation_check.log
done
cat gpurun_out/mutation_check.log
