import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """The CPU oracle is test infrastructure: compile it if the prebuilt file is not there.
    (libphmm_cuda.so itself is built by __graft_entry__.build(); tests never build it silently.)"""
    import oracle
    oracle.build()


def load_golden(name):
    from genomicsbench_palisade_b200.batches import FlatBatches
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    fb = FlatBatches(z["batch_num_reads"], z["batch_num_haps"], z["read_len"], z["read_off"], z["rs"],
                     z["q"], z["i"], z["d"], z["c"], z["hap_len"], z["hap_off"], z["hap"])
    return fb, {k: z[k] for k in ("out", "raw32", "raw64", "rescued", "out_f32only", "out_f64only")}


def same_or_close(a, b, rtol):
    """Relative comparison that treats equal infinities (log10 of an underflowed sum) as equal."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    eq = (a == b)
    with np.errstate(invalid="ignore", divide="ignore"):
        rel = np.abs(a - b) / np.abs(b)
    return bool(np.all(eq | (rel <= rtol)))


def max_rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        rel = np.where(a == b, 0.0, np.abs(a - b) / np.abs(b))
    return float(np.nanmax(rel)) if rel.size else 0.0


GOLDEN_NAMES = ["kat", "tiny", "edge", "c1_small", "c2_large", "c3_regions", "c4_stress"]
