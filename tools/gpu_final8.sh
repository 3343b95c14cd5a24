set -x
N=${1:-8}
mkdir -p gpurun_out
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $N --steps 20 --warmup 5 ) > gpurun_out/r2_final_bench_n$N.json 2> gpurun_out/r2_final_bench_n$N.err
tail -4 gpurun_out/r2_final_bench_n$N.err
python - $N <<'PY'
import json,sys
N=sys.argv[1]
d=json.loads([l for l in open('gpurun_out/r2_final_bench_n%s.json'%N) if l.startswith('{')][-1])
print("N", d["n_gpus"], "value", d["value"], "ms", d["ms_per_step"], "imbalance", d["imbalance"], "e2e", d["e2e"]["value"], d["e2e"]["matches_resident_run"], "parity", d["parity"]["ok"], d["parity"]["pairs"])
for k,b in d.get("configs",{}).items():
    print(k, "value %.0f e2e %.0f frac %.3f imb %.4f parity %s e2e_ok %s"%(b["value"], b["e2e"]["value"], b["frac"], b["imbalance"], b["parity"]["ok"], b["e2e"]["matches_resident_run"]), b.get("speedup_vs_n1"), b.get("n1_ms_per_step"))
PY
