"""Shared helpers for the parity tests (CPU-side; the oracle is the checker)."""
import glob
import os

import numpy as np

from oracle import oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN, "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


def golden_stream(g):
    return O.Stream(g["pos"], g["code"], g["table"], int(g["T"]), g["block_bounds"], g["block_index_bounds"],
                    int(g["s_rows"]), int(g["s_cols"]))


def bits(x):
    return np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)


def assert_bits_equal(x, y, what=""):
    bx, by = bits(x), bits(y)
    assert bx.shape == by.shape, f"{what}: shape {bx.shape} vs {by.shape}"
    if not np.array_equal(bx, by):
        bad = np.flatnonzero(bx.reshape(-1) != by.reshape(-1))
        raise AssertionError(f"{what}: {bad.size} of {bx.size} elements differ, first at {bad[0]}: "
                             f"{np.asarray(x).reshape(-1)[bad[0]]!r} vs {np.asarray(y).reshape(-1)[bad[0]]!r}")
