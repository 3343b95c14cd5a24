mkdir -p gpurun_out
run() { tag=$1; shift
  env "$@" python bench.py --no-cpu --steps 10 --configs c1,c3 > gpurun_out/r2_e2e_$tag.json 2> gpurun_out/r2_e2e_$tag.err || tail -3 gpurun_out/r2_e2e_$tag.err
  python - $tag <<'PY'
import json,sys
d=json.loads([l for l in open('gpurun_out/r2_e2e_%s.json'%sys.argv[1]) if l.startswith('{')][-1])
def f(b): return "res %.0f e2e %.0f (%.3f) pinned %.0f (%.3f) ok %s"%(b["value"], b["e2e"]["value"], b["e2e"]["value"]/b["value"], (b["e2e"].get("pinned") or {}).get("value", b["e2e"].get("pinned_value",0)), (b["e2e"].get("pinned") or {}).get("value", b["e2e"].get("pinned_value",0))/b["value"], b["e2e"]["matches_resident_run"])
print(sys.argv[1], "c2", f(d))
for k,b in d["configs"].items(): print(sys.argv[1], k, f(b))
PY
}
run p1 PHMM_PLANNERS=1
run p3 PHMM_PLANNERS=3
run p3g2 PHMM_PLANNERS=3 PHMM_CHUNK_SCHED=64,2,3,4
run p4f32 PHMM_PLANNERS=4 PHMM_CHUNK_SCHED=32,4,3,4
nproc
