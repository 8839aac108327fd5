#!/bin/bash
# compute-sanitizer over the kernel tests (tiny shapes) and one eager step with side streams on (SURVEY section 5:
# "memcheck / racecheck on every kernel").  Run on a B200:  bash tools/sanitize.sh > gpurun_out/r02_sanitizer.txt 2>&1
# memcheck: out-of-bounds / misaligned global + shared accesses, leaks of device memory are not checked (torch caches).
# racecheck: shared-memory hazards inside a CTA (the mbarrier / DSMEM protocols of lstm.cu, the attention tiles).
# synccheck: invalid barrier usage (named barriers of the fused-norm epilogue, cluster barriers).
cd "$(dirname "$0")/.."
SAN=/usr/local/cuda/bin/compute-sanitizer
small="hiddenEncoder=64 hiddenGar=64 nPredicts=4 negativeSamplingExt=16 sizeWindow=5120 batchSizeGPU=4"
mid="nPredicts=4 negativeSamplingExt=32 sizeWindow=5120 batchSizeGPU=4"
run() {
  echo "=== $*"
  timeout 1500 "$@" 2>&1 | grep -E "ERROR SUMMARY|RACECHECK SUMMARY|SYNCCHECK|Error|error|hazard|loss \[|passed|failed" | sort | uniq -c | head -20
}
for tool in memcheck racecheck synccheck; do
  run $SAN --tool $tool --print-limit 5 python tools/step_once.py $small
  run $SAN --tool $tool --print-limit 5 python tools/step_once.py $mid
done
run $SAN --tool memcheck --print-limit 5 python tools/step_once.py $small arMode=transformer
run $SAN --tool memcheck --print-limit 5 python tools/step_once.py $small arMode=GRU nLevelsGRU=2 rnnMode=linear
run $SAN --tool memcheck --print-limit 5 python -m pytest tests/test_kernels_gpu.py -q -x -m gpu -k "not two_gpus and not 20480 and not 2048"
