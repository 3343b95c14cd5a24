set -x
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x ) > gpurun_out/r2c_pytest_gpu.log 2>&1; tail -5 gpurun_out/r2c_pytest_gpu.log
python bench.py --workload c4 --no-cpu --steps 5 > gpurun_out/r2c_c4.json 2> gpurun_out/r2c_c4.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2c_c4.json').read().strip().split('\n')[-1])
print("C4 value", d["value"], "ms32", d["roofline"]["ms_fp32"], "ms64", d["roofline"]["ms_fp64"], d["e2e"]["matches_resident_run"])
PY
python bench.py --workload c2 --configs none --no-cpu --steps 5 > gpurun_out/r2c_c2.json 2> gpurun_out/r2c_c2.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2c_c2.json').read().strip().split('\n')[-1])
print("C2 value", d["value"], "ms32", d["roofline"]["ms_fp32"], "ms64", d["roofline"]["ms_fp64"], d["e2e"]["matches_resident_run"])
PY
