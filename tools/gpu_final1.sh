set -x
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q ) > gpurun_out/r2_final_pytest_gpu.log 2>&1; tail -3 gpurun_out/r2_final_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_final_smoke.log 2>&1; tail -3 gpurun_out/r2_final_smoke.log
( time python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 ) > gpurun_out/r2_final_bench_reference_n1.json 2> gpurun_out/r2_final_bench_n1.err
( time python bench.py --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/r2_final_bench_n1.json 2>> gpurun_out/r2_final_bench_n1.err; tail -4 gpurun_out/r2_final_bench_n1.err
