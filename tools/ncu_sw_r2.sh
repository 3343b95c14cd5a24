#!/bin/bash
# ncu --set full of every sw_align_kernel launch of one warm step of tools/sw_bench.py -> DRAM traffic per step.
set -e
OUT=${1:-gpurun_out}
TAG=${2:-r2_sw}
CMD="python tools/sw_bench.py --steps 1 --warmup 1 --no-cpu --no-e2e"
ncu --set full --clock-control none -k sw_align_kernel --launch-skip 16 -c 16 -f -o /tmp/sw $CMD > $OUT/${TAG}_ncu.log 2>&1 || true
ncu -i /tmp/sw.ncu-rep --page raw --csv > /tmp/sw_raw.csv
python tools/ncu_pick.py /tmp/sw_raw.csv > $OUT/${TAG}_ncu_full_metrics.txt
python - "$OUT" "$TAG" <<'PY'
import csv, json, sys, time
out, tag = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open("/tmp/sw_raw.csv")))
hdr, units = rows[0], rows[1]
def val(r, i):
    v = float(r[i].replace(",", ""))
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(units[i].lower(), 1)
rd, wr, du = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
R = sum(val(r, rd) for r in rows[2:]); W = sum(val(r, wr) for r in rows[2:])
dur_unit = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "msecond": 1.0, "ms": 1.0}.get(units[du].lower(), 1e-6)
D = sum(float(r[du].replace(",", "")) for r in rows[2:]) * dur_unit
d = {"sw": {"kernel": "sw_align_kernel", "launches": len(rows) - 2, "dram_read_bytes": R, "dram_write_bytes": W,
            "traffic_bytes_per_step": R + W, "sum_duration_ms": D,
            "source": "ncu --set full --clock-control none -k sw_align_kernel, one warm step of tools/sw_bench.py (100 000 pairs), captured " + time.strftime("%Y-%m-%d")}}
json.dump(d, open(out + "/" + tag + "_dram_traffic.json", "w"), indent=1)
print(json.dumps(d))
PY
