mkdir -p gpurun_out
for by in time cells; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 10 --warmup 3 --workload skew --shard-by $by --no-cpu > gpurun_out/r2_skew_n2_by_$by.json 2> gpurun_out/r2_skew.err || tail -5 gpurun_out/r2_skew.err
python - $by <<'PY'
import json,sys
d=json.loads([l for l in open('gpurun_out/r2_skew_n2_by_%s.json'%sys.argv[1]) if l.startswith('{')][-1])
print("by", sys.argv[1], "value %.0f ms %.2f imbalance %s e2e %.0f rescue frac %.3f"%(d["value"], d["ms_per_step"], d["imbalance"], d["e2e"]["value"], d["rescue"]["frac"]))
PY
done
