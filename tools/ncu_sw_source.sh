#!/bin/bash
# Profile artefacts of the Smith-Waterman row (run from the repo root on the GPU box):
#   $OUT/sw_bench.json         tools/sw_bench.py, no profiler
#   $OUT/sw_launches.csv       ncu launch list (gpu__time_duration.sum) of one step
#   $OUT/sw_ncu_full.txt       selected metrics of an ncu --set full capture of the launches of one step
#   $OUT/sw_source.csv         SASS source page (per-instruction stalls) of the longest launch
set -e
OUT=${1:-gpurun_out}
python tools/sw_bench.py > $OUT/sw_bench.json 2> $OUT/sw_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -k sw_align_kernel -c 40 --csv --log-file $OUT/sw_launches.csv \
    python tools/sw_bench.py --steps 1 --warmup 0 --no-cpu --no-e2e > /dev/null 2>&1
ncu --set full --import-source on --clock-control none -k sw_align_kernel -c 24 -f -o /tmp/swk \
    python tools/sw_bench.py --steps 1 --warmup 0 --no-cpu --no-e2e > $OUT/sw_ncu.log 2>&1
ncu -i /tmp/swk.ncu-rep --page raw --csv > /tmp/sw_raw.csv
python tools/ncu_pick.py /tmp/sw_raw.csv > $OUT/sw_ncu_full.txt
python - "$OUT" <<'PY'
import csv, sys, subprocess
out = sys.argv[1]
rows = list(csv.reader(open("/tmp/sw_raw.csv")))
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
