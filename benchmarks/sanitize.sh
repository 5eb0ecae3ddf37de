#!/usr/bin/env bash
# compute-sanitizer over the GPU tests at tiny shapes.  ONE tool per invocation (the profiling guide of this pool:
# several tools in one call have left GPUs unusable), on a program that already passed without the tool:
#
#   benchmarks/sanitize.sh memcheck   [pytest -k expression]
#   benchmarks/sanitize.sh racecheck  [pytest -k expression]
#   benchmarks/sanitize.sh synccheck  [pytest -k expression]
#
# Output: gpurun_out/sanitize_<tool>.log (copy the summary line to profiles/ when it is to be kept).
# The multi-rank flag protocol is covered by the single-GPU emulation tests (tests/test_gpu_fabric.py: all ranks
# as co-resident slices of one cooperative grid); real multi-GPU runs are never wrapped in a tool.
set -uo pipefail
TOOL=${1:-memcheck}
EXPR=${2:-"engine_step_matches_autograd or fused_grad_aggregation or ctm_gemms or theta_mc or csr_to_csc"}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export GFEDNTM_NO_GRAPHS=1            # eager launches: every kernel is visible to the tool
PYT="python -m pytest tests/test_gpu_engine.py tests/test_gpu_fabric.py tests/test_gpu_ctm_gemm.py tests/test_gpu_decoder.py -x -q -m gpu -k"
# 1) the same command must pass without the tool first
$PYT "$EXPR" > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed, not sanitizing"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
# 2) one tool
compute-sanitizer --tool "$TOOL" --error-exitcode 3 --launch-timeout 600 $PYT "$EXPR" > "gpurun_out/sanitize_${TOOL}.log" 2>&1
rc=$?
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" "gpurun_out/sanitize_${TOOL}.log" | tail -5
exit $rc
