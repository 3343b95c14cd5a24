mkdir -p gpurun_out
python -m pytest tests/test_sw.py -m gpu -q -x 2>&1 | tail -3
python tools/sw_bench.py --no-cpu > gpurun_out/r2_sw_bench.json 2> gpurun_out/r2_sw_bench.err || tail -5 gpurun_out/r2_sw_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_sw_bench.json').read().strip().split('\n')[-1])
print("SW value %.0f GCUPS (%.2f ms) e2e %.0f (%.3f) same %s launches %d"%(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["value"]/d["value"], d["e2e"]["matches_resident_run"], d["gpu_launches"]))
PY
bash tools/ncu_sw_r2.sh gpurun_out r2_sw 2>&1 | tail -2
