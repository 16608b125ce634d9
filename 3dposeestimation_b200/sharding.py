"""Host-side arithmetic of the row-sharded template DB (new - the reference is single process).

The DB of T rows is cut into `world` contiguous row ranges; rank r holds rows [lo, hi) and reports
global row indices (local index + lo).  Every rank scans its shard for ALL queries, the per-shard
(distance, row) lists are all-gathered and merged with the matcher's own ordering: ascending distance,
ties -> lower global row (what a stable single-shard scan returns, reference
src/networkevaluator.cpp:197,267 via OpenCV BFMatcher).  The CUDA path does the scan and the merge in
kernels (csrc/knn*.cu, csrc/comm.cu); this module only holds the index arithmetic that bench.py, the
tests and the C++ host mirror share, plus a numpy statement of the merge rule used as a checker.
"""
import numpy as np


def shard_rows(rank: int, world: int, total: int):
    """rows [lo, hi) of rank `rank`: balanced to within one row, contiguous, covering [0, total)"""
    if not (0 <= rank < world) or total < 0:
        raise ValueError("bad rank / world / total")
    return rank * total // world, (rank + 1) * total // world


def merge_rule(idx_lists, dist_lists, k: int):
    """checker for the merge kernel: idx/dist lists are [world][nq][k'] with -1 / inf padding"""
    idx = np.concatenate(idx_lists, axis=1)
    dist = np.concatenate(dist_lists, axis=1).astype(np.float64)
    key_idx = np.where(idx < 0, np.iinfo(np.int64).max, idx.astype(np.int64))
    order = np.lexsort((key_idx, dist), axis=1)[:, :k]
    out_i = np.take_along_axis(idx, order, 1)
    out_d = np.take_along_axis(np.concatenate(dist_lists, axis=1), order, 1)
    return out_i, out_d


def merge_histograms(hists, errs, bins):
    """checker for pe_histogram_sharded_dev: per-rank computeHistogram outputs (n_bins + 2 floats each: normalised bins,
    mean, median) + per-rank per-query errors (-1 = no same-class neighbour) -> the histogram of the whole query set, as
    the reference's single loop would produce it (src/networkevaluator.cpp:337-350): counts add, the mean is the error
    sum over ALL queries, the median is element (n_valid - 1) / 2 of the sorted valid errors"""
    bins = np.asarray(bins, np.float32)
    nb = bins.size
    counts = np.zeros(nb, np.float64)
    total_sum, nq = 0.0, 0
    for h, e in zip(hists, errs):
        e = np.asarray(e, np.float32)
        valid = e >= 0
        tot_r = float((~valid).sum() + (valid & (e < bins[-1])).sum())     # what the rank divided its bins by
        counts += np.round(np.asarray(h[:nb], np.float64) * tot_r)
        total_sum += float(e[valid].astype(np.float64).sum())
        nq += e.size
    out = np.empty(nb + 2, np.float32)
    out[:nb] = (counts.astype(np.float32) / np.float32(counts[0] + counts[nb - 1]))
    out[nb] = np.float32(total_sum / nq) if nq else np.float32(np.nan)
    v = np.sort(np.concatenate([np.asarray(e, np.float32)[np.asarray(e) >= 0] for e in errs]))
    out[nb + 1] = v[(v.size - 1) // 2] if v.size else np.float32(np.nan)
    return out
