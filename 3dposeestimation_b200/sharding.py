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
