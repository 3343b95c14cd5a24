#!/bin/bash
# ncu --set full of the FP64 rescue kernels of the C4 workload; exports the raw page of all captured
# launches and the SASS source page (per-instruction stalls) of the longest FP64 launch.
set -e
OUT=${1:-gpurun_out}
ncu --set full --import-source on --clock-control none -k phmm_forward_kernel -c 12 -f -o /tmp/f64 \
    python bench.py --workload c4 --steps 1 --warmup 1 --no-cpu > $OUT/f64_ncu.log 2>&1
ncu -i /tmp/f64.ncu-rep --page raw --csv > $OUT/f64_raw.csv
python - "$OUT" <<'PY'
import csv, sys, subprocess
out = sys.argv[1]
rows = list(csv.reader(open(out + "/f64_raw.csv")))
hdr = rows[0]
ki, di = hdr.index("Kernel Name"), hdr.index("gpu__time_duration.sum")
best, bid = -1.0, None
for n, r in enumerate(rows[2:]):
    if "double" in r[ki] or "<double" in r[ki] or "Id" in r[ki]:
        d = float(r[di].replace(",", ""))
        if d > best: best, bid = d, n
print("longest FP64 launch index", bid, best, rows[2 + bid][ki])
src = subprocess.run(["ncu", "-i", "/tmp/f64.ncu-rep", "--page", "source", "--csv", "--print-source", "sass",
                      "--launch-skip", str(bid), "--launch-count", "1"], capture_output=True, text=True).stdout
open(out + "/f64_source.csv", "w").write(src)
PY
