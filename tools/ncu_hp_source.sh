#!/bin/bash
# ncu --set full of the fast FP32 kernels (phmm_forward_hp_kernel) of the default bench.py workload; exports the
# raw page of the captured launches and the SASS source page (per-instruction stalls) of the longest one.
set -e
OUT=${1:-gpurun_out}
ncu --set full --import-source on --clock-control none -k phmm_forward_hp_kernel -c 12 -f -o /tmp/hp \
    python bench.py --steps 1 --warmup 1 --no-cpu > $OUT/hp_ncu.log 2>&1
ncu -i /tmp/hp.ncu-rep --page raw --csv > /tmp/hp_raw.csv
python tools/ncu_pick.py /tmp/hp_raw.csv > $OUT/hp_ncu_full.txt
python - "$OUT" <<'PY'
import csv, sys, subprocess
out = sys.argv[1]
rows = list(csv.reader(open("/tmp/hp_raw.csv")))
hdr = rows[0]
ki, di = hdr.index("Kernel Name"), hdr.index("gpu__time_duration.sum")
best, bid = -1.0, None
for n, r in enumerate(rows[2:]):
    d = float(r[di].replace(",", ""))
    if d > best: best, bid = d, n
print("longest launch index", bid, best, rows[2 + bid][ki])
src = subprocess.run(["ncu", "-i", "/tmp/hp.ncu-rep", "--page", "source", "--csv", "--print-source", "sass",
                      "--launch-skip", str(bid), "--launch-count", "1"], capture_output=True, text=True).stdout
open(out + "/hp_source.csv", "w").write(src)
PY
