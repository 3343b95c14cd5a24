#!/bin/bash
# ncu --set full of the Smith-Waterman class kernels of tools/sw_bench.py; exports the raw page of the captured
# launches and the SASS source page (per-instruction stalls) of the longest one.
set -e
OUT=${1:-gpurun_out}
ncu --set full --import-source on --clock-control none -k sw_align_kernel -c 12 -f -o /tmp/swk \
    python tools/sw_bench.py --steps 1 --warmup 0 --no-cpu --no-e2e > $OUT/sw_ncu.log 2>&1
ncu -i /tmp/swk.ncu-rep --page raw --csv > $OUT/sw_raw.csv
python - "$OUT" <<'PY'
import csv, sys, subprocess
out = sys.argv[1]
rows = list(csv.reader(open(out + "/sw_raw.csv")))
hdr = rows[0]
ki, di = hdr.index("Kernel Name"), hdr.index("gpu__time_duration.sum")
best, bid = -1.0, None
for n, r in enumerate(rows[2:]):
    d = float(r[di].replace(",", ""))
    if d > best: best, bid = d, n
print("longest launch index", bid, best, rows[2 + bid][ki])
src = subprocess.run(["ncu", "-i", "/tmp/swk.ncu-rep", "--page", "source", "--csv", "--print-source", "sass",
                      "--launch-skip", str(bid), "--launch-count", "1"], capture_output=True, text=True).stdout
open(out + "/sw_source.csv", "w").write(src)
PY
