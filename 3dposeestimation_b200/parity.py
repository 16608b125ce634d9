"""Acceptance metrics of BASELINE.json / SURVEY 8(c), shared by tests/, __graft_entry__.smoke() and bench.py.
Pure numpy; nothing here computes a reference result - the callers bring the oracle's output.

Two readings of "descriptors within 1e-4 relative (fp32)" are reported side by side:

  row-scaled        max |a-b| / max|b_row|  < 1e-4      the whole descriptor is the unit the matcher consumes
  element-relative  |a-b| <= 1e-4 * max(|a|, |b|, eps_row), eps_row = 1e-4 * max|b_row|     (SURVEY 8c verbatim)

The second cannot be met by EVERY element of any fp32 implementation that sums in another order than Caffe (an
element 1e3 times smaller than the row max carries the row's ~1e-7 rounding noise as a 1e-4 relative error), so it is
reported as the fraction of elements that meet it; BASELINE.json's target is >= 90 % agreement.
"""
import numpy as np

DESC_TOL = 1e-4


def rowscale_err(a, b):
    """max over elements of |a-b| / (max-abs of the reference row)"""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    scale = np.abs(b).max(axis=-1, keepdims=True) + 1e-30
    return float((np.abs(a - b) / scale).max())


def elementrel(a, b, tol=DESC_TOL):
    """per element |a-b| <= tol * max(|a|, |b|, eps_row).  Returns (fraction of elements that pass,
    fraction of rows whose every element passes, worst ratio |a-b| / max(|a|,|b|,eps_row))"""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    if a.size == 0:
        return 1.0, 1.0, 0.0
    eps_row = tol * np.abs(b).max(axis=-1, keepdims=True)
    den = np.maximum(np.maximum(np.abs(a), np.abs(b)), eps_row) + 1e-300
    ratio = np.abs(a - b) / den
    ok = ratio <= tol
    return float(ok.mean()), float(ok.all(axis=-1).mean()), float(ratio.max())


def knn_agreement(idx, oracle_idx, oracle_dist, oracle_gap, tol=DESC_TOL):
    """nearest-template indices 'bit-exact except where the top-2 (k / k+1, or an inner) distance gap is below tol':
    returns (fraction of rows identical, fraction of rows that differ WITHOUT a near-tie to excuse them)"""
    idx, oi, od = np.asarray(idx), np.asarray(oracle_idx), np.asarray(oracle_dist, np.float64)
    if idx.shape[0] == 0:
        return 1.0, 0.0
    bad = (idx != oi).any(axis=1)
    inner = (np.diff(od, axis=1) / np.maximum(od[:, 1:], 1e-30)) if od.shape[1] > 1 else np.full((len(od), 1), np.inf)
    near = (np.asarray(oracle_gap) < tol) | (np.nan_to_num(inner, nan=np.inf).min(axis=1) < tol)
    return float(1.0 - bad.mean()), float((bad & ~near).mean())
