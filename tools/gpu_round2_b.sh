set -x
mkdir -p gpurun_out
tools/ubench6 > gpurun_out/r2_ubench6_fp64_issue.log 2>&1; cat gpurun_out/r2_ubench6_fp64_issue.log
( time python -m pytest tests -m gpu -q ) > gpurun_out/r2b_pytest_gpu.log 2>&1; tail -5 gpurun_out/r2b_pytest_gpu.log
