#!/bin/bash
# compute-sanitizer pass over the small-scene GPU tests (the `debug` contract of the reference is
# notes/rasterizer_note.h:44-53).  Run on the GPU box:  gpurun -- 'tools/sanitize.sh r2'
#   -> gpurun_out/sanitizer_<tag>_{memcheck,racecheck,initcheck,synccheck}.txt  (copy the summaries into profiles/)
# The test selection keeps every kernel of the library on the path (preprocess, tile binning, onesweep sort, both
# compositing kernels, preprocess backward, loss, Adam) at sizes the tools finish in minutes.
set -u
TAG=${1:-r2}
SEL=${SANITIZE_SEL:-"tests/test_gpu_parity.py tests/test_gpu_tile_binning.py tests/test_gpu_api.py tests/test_gpu_training.py tests/test_gpu_sort.py"}
KEXPR=${SANITIZE_K:-"not c2 and not full_size and not 20000 and not subprocess"}
mkdir -p gpurun_out
export GSR_SANITIZE=1          # tests shrink their largest cases under the tools
export PYTORCH_NO_CUDA_MEMORY_CACHING=1   # every torch.empty is its own cudaMalloc: out-of-bounds accesses cannot hide in the cache
for TOOL in memcheck racecheck initcheck synccheck; do
  OUT=gpurun_out/sanitizer_${TAG}_${TOOL}.txt
  EXTRA=""
  [ "$TOOL" = "initcheck" ] && EXTRA="--track-unused-memory no"
  [ "$TOOL" = "memcheck" ] && EXTRA="--leak-check no"
  echo "== $TOOL" | tee $OUT
  timeout ${SANITIZE_TIMEOUT:-900} compute-sanitizer --tool $TOOL $EXTRA --print-limit 20 --error-exitcode 0 \
      --kernel-name kns=gsr --log-file gpurun_out/sanitizer_${TAG}_${TOOL}.log \
      python -m pytest $SEL -q -x -m gpu -k "$KEXPR" -p no:cacheprovider 2>&1 | tail -4 | tee -a $OUT
  echo "-- tool log (errors / summary)" >> $OUT
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error:|Warning:|Race reported|Uninitialized|Invalid|hazard" gpurun_out/sanitizer_${TAG}_${TOOL}.log | sort | uniq -c | sort -rn | head -40 >> $OUT
  tail -3 gpurun_out/sanitizer_${TAG}_${TOOL}.log >> $OUT
done
