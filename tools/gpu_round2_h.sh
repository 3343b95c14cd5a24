mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py --no-cpu --steps 10 --configs c1,c3,c4 > gpurun_out/r2h_e2e.json 2> gpurun_out/r2h_e2e.err || tail -3 gpurun_out/r2h_e2e.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2h_e2e.json') if l.startswith('{')][-1])
def f(b): return "res %.0f e2e %.0f (%.3f) pinned %.0f (%.3f) ok %s"%(b["value"], b["e2e"]["value"], b["e2e"]["value"]/b["value"], (b["e2e"].get("pinned") or {}).get("value", b["e2e"].get("pinned_value",0)), (b["e2e"].get("pinned") or {}).get("value", b["e2e"].get("pinned_value",0))/b["value"], b["e2e"]["matches_resident_run"])
print("c2", f(d))
for k,b in d["configs"].items(): print(k, f(b))
PY
python tools/gatk_bench.py --kernels exact --no-cpu 2>/dev/null | python -c "
import json,sys
g=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('gatk exact', [(x['regions_per_call'], round(x['ms_per_call_median'],3), round(x['gcups'])) for x in g['groups']])"
python tools/gatk_bench.py --kernels fast --no-cpu 2>/dev/null | python -c "
import json,sys
g=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('gatk fast', [(x['regions_per_call'], round(x['ms_per_call_median'],3), round(x['gcups'])) for x in g['groups']])"
python tools/legacy_driver_bench.py --threads 16 --no-ref 2>&1 >/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l); print('legacy', r['workload'], 'fast', r['PHMM_FAST'], 'coalesce', r['coalesce'], round(r['gcups']))"
