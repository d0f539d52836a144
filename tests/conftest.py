import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "convex-optimization_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, "ref_%s.npz" % name), allow_pickle=False)
    rec = {k: z[k] for k in z.files}
    opts = {}
    for k, v in zip(rec["opts_keys"].tolist(), rec["opts_vals"].tolist()):
        try:
            opts[k] = int(v)
        except ValueError:
            try:
                opts[k] = float(v)
            except ValueError:
                opts[k] = v
    rec["opts"] = opts
    for k in ("seed", "m", "n", "l"):
        rec[k] = int(rec[k])
    rec["mu"] = float(rec["mu"])
    rec["in_dtype"] = str(rec["in_dtype"])
    return rec


def golden_inputs(rec):
    """Regenerate the inputs of a golden case (same restated gen_data the generator used)."""
    from oracle.glprox_oracle import gen_data
    A, b, x0, u, mu = gen_data(rec["seed"], rec["m"], rec["n"], rec["l"])
    if rec["in_dtype"] == "f4":
        A, b, x0 = (z.astype(np.float32).astype(np.float64) for z in (A, b, x0))
    return A, b, x0, u, mu


@pytest.fixture
def golden():
    return load_golden
