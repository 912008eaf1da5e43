#!/bin/bash
# compute-sanitizer over the kernel numerics tests: ONE tool per gpurun call (B200_PROFILING.md).
#   gpurun -- 'scripts/sanitize.sh memcheck'      (or racecheck / synccheck / initcheck)
# Run the plain tests first; the sanitizer only runs if they exit 0.
set -e
TOOL=${1:-memcheck}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout 300 > gpurun_out/sanitize_plain.log 2>&1
compute-sanitizer --tool "$TOOL" --error-exitcode 9 --log-file gpurun_out/sanitize_$TOOL.log \
    python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout 1200 -k "gemm or conv2 or head or adam or push" \
    > gpurun_out/sanitize_$TOOL.out 2>&1
tail -5 gpurun_out/sanitize_$TOOL.out
grep -E "ERROR SUMMARY|RACECHECK SUMMARY" gpurun_out/sanitize_$TOOL.log | tail -3
