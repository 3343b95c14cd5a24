set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv
nproc; free -g | head -2
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2a_pytest_gpu.log 2>&1; tail -5 gpurun_out/r2a_pytest_gpu.log
( time python bench.py --steps 5 --warmup 3 ) > gpurun_out/r2a_bench_n1.json 2> gpurun_out/r2a_bench_n1.err; tail -c 3000 gpurun_out/r2a_bench_n1.json; tail -5 gpurun_out/r2a_bench_n1.err
