import sys, time, numpy as np
sys.path.insert(0, '.')
import bench
from genomicsbench_palisade_b200 import phmm
name = sys.argv[1]
fb = bench.make_workload(name, 1)
eng = phmm.PairHMM(fast=True, fast64=True)
o = np.empty(fb.n_pairs)
for k in range(3):
    t = time.perf_counter(); eng.compute(fb, out=o); print("TOTAL %s %.2f ms" % (name, (time.perf_counter() - t) * 1e3), file=sys.stderr)
