"""Helpers around the golden fixtures in tests/golden/ (written by oracle/gen_golden.py from the
unmodified reference) and the parity metric of the reference's own tests."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


class Stages:
    """Sequential log: stage 0 holds every watched array, later stages the arrays each call changed."""

    def __init__(self, fname):
        self.z = np.load(os.path.join(GOLDEN, fname))
        self.labels = json.loads(bytes(self.z["__labels__"]).decode())
        self.index = {}
        for n, (label, _) in enumerate(self.labels):
            self.index.setdefault(label, n)

    def state_at(self, n):
        """All watched arrays after stage n (inclusive)."""
        out = {}
        for m in range(n + 1):
            for name in self.labels[m][1]:
                out[name] = self.z["%03d:%s" % (m, name)]
        return out

    def before(self, label):
        return self.state_at(self.index["pre:" + label])

    def after(self, label):
        return self.state_at(self.index[label])

    def changed_by(self, label):
        return list(self.labels[self.index[label]][1])


def load_grid(fname):
    z = np.load(os.path.join(GOLDEN, fname))
    return {k: z[k] for k in z.files}


def rel_err(a, b):
    """Reference metric (NDSL/ndsl/testing/comparison.py:6-68): 2|a-b|/(|a|+|b|), 0 where both are 0."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    denom = np.abs(a) + np.abs(b)
    with np.errstate(invalid="ignore", divide="ignore"):
        r = 2.0 * np.abs(a - b) / denom
    r[denom == 0] = 0.0
    r[np.isnan(a) & np.isnan(b)] = 0.0
    return r


def report(name, got, want, near_zero=0.0):
    """max relative error (ignoring |values| below near_zero), number of non-bit-identical elements"""
    r = rel_err(got, want)
    if near_zero:
        r[(np.abs(got) < near_zero) & (np.abs(want) < near_zero)] = 0.0
    got, want = np.asarray(got), np.asarray(want)
    both_nan = np.isnan(got) & np.isnan(want)
    nbad = int(np.count_nonzero(got != want) - np.count_nonzero(both_nan))
    # A NaN / Inf on one side only is a failure, not a zero: the reference metric lets NaN pass only against NaN
    # (NDSL/ndsl/testing/comparison.py:6-68), and 2|inf - x| / (inf + |x|) is itself NaN.
    nonfinite = (~np.isfinite(got) | ~np.isfinite(want)) & ~both_nan & (got != want)
    if nonfinite.any():
        return float("inf"), nbad
    return float(np.nanmax(r)) if r.size else 0.0, nbad


def passes(err, nbad, tol):
    """tol == 0.0 is the bit-exact bar: no element may differ at all (the metric alone would let a sign of zero or a
    NaN-vs-NaN pair through, which is what the reference accepts too); otherwise the metric decides."""
    return (nbad == 0 and err == 0.0) if tol == 0.0 else err <= tol
