mkdir -p gpurun_out
for sched in default 0.08,0.32,0.4,0.2 0.06,0.47,0.47 0.1,0.9 0.05,0.25,0.5,0.2 1.0; do
  if [ "$sched" != default ]; then export SW_CHUNK_SCHED=$sched; fi
  python tools/sw_bench.py --no-cpu --steps 8 > gpurun_out/r2_sw_bench_$sched.json 2> gpurun_out/r2_sw_bench.err || tail -5 gpurun_out/r2_sw_bench.err
  python - $sched <<'PY'
import json,sys
d=json.loads(open('gpurun_out/r2_sw_bench_%s.json'%sys.argv[1]).read().strip().split('\n')[-1])
print(sys.argv[1], "SW value %.0f GCUPS (%.2f ms) e2e %.0f (%.3f, %.2f ms) same %s"%(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["value"]/d["value"], d["e2e"]["ms_per_step"], d["e2e"]["matches_resident_run"]))
PY
done
