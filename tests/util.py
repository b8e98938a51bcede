"""Shared helpers for the parity tests."""
import numpy as np

from oracle import FIELD_NAMES


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def assert_bitwise(got: dict, want: dict, names=FIELD_NAMES, tag=""):
    bad = []
    for n in names:
        if not np.array_equal(bits(got[n]), bits(want[n])):
            d = np.abs(got[n].astype(np.float64) - want[n].astype(np.float64))
            where = np.unravel_index(np.argmax(d), d.shape)
            bad.append(f"{n}: {np.count_nonzero(bits(got[n]) != bits(want[n]))} cells differ, "
                       f"max abs {d.max():.3e} at (z,y,x)={where}")
    assert not bad, f"{tag} not bit-identical:\n  " + "\n  ".join(bad)


def rel_l2(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    den = np.sqrt((b * b).sum())
    return float(np.sqrt(((a - b) ** 2).sum()) / den) if den > 0 else float(np.sqrt(((a - b) ** 2).sum()))
