set -x
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x ) > gpurun_out/r2e_pytest_gpu.log 2>&1; tail -5 gpurun_out/r2e_pytest_gpu.log
python tools/legacy_driver_bench.py --threads 1,16 --no-ref > gpurun_out/r2_legacy_driver_coalesce.json 2> gpurun_out/r2_legacy_driver_coalesce.err; cut -c1-200 gpurun_out/r2_legacy_driver_coalesce.err
