#!/bin/bash
# compute-sanitizer passes over smoke() (exact + fast pair-HMM kernels, FP64 rescue, Smith-Waterman):
# memcheck (global/shared out-of-bounds, misaligned), racecheck (shared-memory staging hazards),
# synccheck (__shfl_*_sync / __syncwarp masks).  SURVEY.md section 5 asks for these three.
# Run on a GPU box: gpurun --timeout 1200 -- 'bash tools/sanitize.sh'
# Output: gpurun_out/sanitize_<tool>.log (the last lines hold the ERROR SUMMARY).
# NOTE: on this round's GPU pool compute-sanitizer is closed by the operators (every pass answers
# "compute-sanitizer is closed on this pool", exit 86); the script is kept for pools where it is open.
# Out-of-bounds writes are looked for with the library's own red zones instead (PHMM_CANARY=1).
set -u
mkdir -p gpurun_out
rc=0
for tool in memcheck racecheck synccheck; do
  timeout "${SANITIZE_TIMEOUT:-360}" compute-sanitizer --tool "$tool" --print-limit 20 \
    --log-file "gpurun_out/sanitize_${tool}.sanlog" \
    python -c "import __graft_entry__ as g; g.smoke()" > "gpurun_out/sanitize_${tool}.log" 2>&1
  r=$?
  echo "exit code $r" >> "gpurun_out/sanitize_${tool}.log"
  echo "== $tool: exit $r"
  tail -n 4 "gpurun_out/sanitize_${tool}.log"
  tail -n 3 "gpurun_out/sanitize_${tool}.sanlog" 2>/dev/null
  [ $r -ne 0 ] && rc=1
done
exit $rc
