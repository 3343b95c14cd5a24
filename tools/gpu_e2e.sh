mkdir -p gpurun_out
for sched in "${@:-default}"; do
  if [ "$sched" != default ]; then export PHMM_CHUNK_SCHED=$sched; fi
  python bench.py --no-cpu --steps 10 > gpurun_out/r2_e2e_$sched.json 2> gpurun_out/r2_e2e_$sched.err || tail -3 gpurun_out/r2_e2e_$sched.err
  python - $sched <<'PY'
import json,sys
d=json.loads([l for l in open('gpurun_out/r2_e2e_%s.json'%sys.argv[1]) if l.startswith('{')][-1])
def f(b): return "res %.0f e2e %.0f (%.3f) pinned %.0f (%.3f) ok %s"%(b["value"], b["e2e"]["value"], b["e2e"]["value"]/b["value"], (b["e2e"].get("pinned") or {}).get("value", b["e2e"].get("pinned_value",0)), (b["e2e"].get("pinned") or {}).get("value", b["e2e"].get("pinned_value",0))/b["value"], b["e2e"]["matches_resident_run"])
print(sys.argv[1], "c2", f(d))
for k,b in d["configs"].items(): print(sys.argv[1], k, f(b))
PY
done
