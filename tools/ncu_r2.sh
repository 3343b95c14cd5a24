#!/bin/bash
# Round-2 ncu evidence for the default bench.py command (C2, one GPU):
#   1. launch list (gpu__time_duration.sum of every kernel of one warm step; cold-cache and serialised: read the shares)
#   2. ncu --set full of the 12 fast FP32 class kernels of one step -> metrics table + DRAM traffic per step
# Run only after the same command has exited 0 without ncu.
set -e
OUT=${1:-gpurun_out}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --configs none"
$CMD > $OUT/r2_ncu_plain_run.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r2_phmm_launches.csv $CMD > $OUT/r2_ncu_launches.log 2>&1
ncu --set full --import-source on --clock-control none -k phmm_forward_hp_kernel -c 12 -f -o /tmp/hp $CMD > $OUT/r2_ncu_full.log 2>&1
ncu -i /tmp/hp.ncu-rep --page raw --csv > /tmp/hp_raw.csv
python tools/ncu_pick.py /tmp/hp_raw.csv > $OUT/r2_phmm_ncu_full_metrics.txt
python - "$OUT" <<'PY'
import csv, json, sys, time
out = sys.argv[1]
rows = list(csv.reader(open("/tmp/hp_raw.csv")))
hdr, units = rows[0], rows[1]
def col(name): return hdr.index(name)
def val(r, i):
    v = float(r[i].replace(",", ""))
    u = units[i].lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
rd, wr, du = col("dram__bytes_read.sum"), col("dram__bytes_write.sum"), col("gpu__time_duration.sum")
R = sum(val(r, rd) for r in rows[2:]); W = sum(val(r, wr) for r in rows[2:])
dur_unit = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "msecond": 1.0, "ms": 1.0}.get(units[du].lower(), 1e-6)
D = sum(float(r[du].replace(",", "")) for r in rows[2:]) * dur_unit
json.dump({"phmm": {"kernel": "phmm_forward_hp_kernel", "launches": len(rows) - 2, "dram_read_bytes": R, "dram_write_bytes": W,
                    "traffic_bytes_per_step": R + W, "sum_duration_ms": D,
                    "build": "round 2 final build (length-paired runs, FP64-first reads; FP32 cell update unchanged since round 1), captured " + time.strftime("%Y-%m-%d"),
                    "source": "ncu --set full --clock-control none -k phmm_forward_hp_kernel -c 12, bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --configs none; profiles/r2_phmm_ncu_full_metrics.txt"}},
          open(out + "/r2_dram_traffic.json", "w"), indent=1)
print(open(out + "/r2_dram_traffic.json").read())
PY
