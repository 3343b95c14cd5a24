import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from conftest import load_golden
from genomicsbench_palisade_b200 import phmm
fb, g = load_golden("edge")
rsl, hpl, _, _ = fb.pairs()
for label, kw in [("both", dict()), ("double-only", dict(precision=2)), ("float-only", dict(precision=1))]:
    eng = phmm.PairHMM(**kw)
    w = eng.workload(fb); st = w.run(); out, raw32, raw64, resc = w.results(raw=True); w.close(); eng.close()
    want = {"both": g["out"], "double-only": g["out_f64only"], "float-only": g["out_f32only"], "scalar": g["out"]}[label]
    with np.errstate(all="ignore"):
        bad = ~((out == want) | (np.abs(out - want) <= 1e-4 * np.abs(want)))
    print(label, "bad", int(bad.sum()), "of", len(out), "launches", st["n_launches"])
    for k in np.nonzero(bad)[0][:12]:
        print("   pair", k, "R", rsl[k], "H", hpl[k], "gpu", out[k], "ref", want[k], "raw32", raw32[k], g["raw32"][k], "raw64", raw64[k], g["raw64"][k], "resc", resc[k], g["rescued"][k])
