mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py --no-cpu --steps 10 --configs c1,c3,c4 > gpurun_out/r2f_e2e.json 2> gpurun_out/r2f_e2e.err || tail -3 gpurun_out/r2f_e2e.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2f_e2e.json') if l.startswith('{')][-1])
def f(b): return "res %.0f e2e %.0f (%.3f) pinned %.0f (%.3f) ok %s"%(b["value"], b["e2e"]["value"], b["e2e"]["value"]/b["value"], (b["e2e"].get("pinned") or {}).get("value", b["e2e"].get("pinned_value",0)), (b["e2e"].get("pinned") or {}).get("value", b["e2e"].get("pinned_value",0))/b["value"], b["e2e"]["matches_resident_run"])
print("c2", f(d))
for k,b in d["configs"].items(): print(k, f(b))
PY
python tools/cli_bench.py --gpus 1 --workloads c2w --loops 20 2>&1 | cut -c1-200 | head -1
