"""Helpers shared by the CPU and GPU tests: the committed golden vectors (tests/golden)."""
import hashlib
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_golden.npz")
CASES = ("testdata", "synthetic_a", "synthetic_c", "attpc_b", "tower800")
ITER_FIELDS = ("votes", "status", "argmax", "a_raw", "b_raw", "a1", "b1", "rc1", "a2", "b2", "rc2",
               "inliers1", "inliers2")


def load():
    return np.load(GOLD)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def case_params(gold, key):
    g, dx, minvotes, nlines = gold[f"{key}_params"]
    return int(g), float(dx), int(minvotes), int(nlines)


def check_trace(gold, key, header, iters):
    """header: dict of scalar fields; iters: list of dicts keyed by ITER_FIELDS."""
    t = f"{key}_trace_"
    for f in ("rc", "n_lines", "n_input", "num_x", "num_b"):
        assert int(header[f]) == int(gold[t + f]), (key, f)
    for f in ("d", "dx", "max_x"):
        assert float(header[f]) == float(gold[t + f]), (key, f)
    assert np.array_equal(np.asarray(header["shift"]), gold[t + "shift"]), key
    assert len(iters) == int(gold[t + "n_iters"]), key
    for k, it in enumerate(iters):
        for f in ITER_FIELDS:
            want = gold[f"{t}it{k}_{f}"]
            got = np.asarray(it[f])
            assert np.array_equal(got, want, equal_nan=True) if got.dtype.kind == "f" else np.array_equal(got, want), (key, k, f)
