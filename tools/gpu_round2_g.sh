mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x -k "golden or random or large_slice" 2>&1 | tail -2
for w in c2 c4 c5; do
python bench.py --workload $w --configs none --no-cpu --steps 5 > gpurun_out/r2g_$w.json 2> gpurun_out/r2g_$w.err || tail -3 gpurun_out/r2g_$w.err
python - $w <<'PY'
import json,sys
d=json.loads([l for l in open('gpurun_out/r2g_%s.json'%sys.argv[1]) if l.startswith('{')][-1])
print(sys.argv[1], "value %.0f ms32 %.2f ms64 %.2f e2e %.0f ok %s rescued %d launches %d"%(d["value"], d["roofline"]["ms_fp32"], d["roofline"]["ms_fp64"], d["e2e"]["value"], d["e2e"]["matches_resident_run"], d["rescue"]["pairs"], d["gpu_launches"]))
PY
done
