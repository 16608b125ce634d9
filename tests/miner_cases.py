"""Seeded sample sets for the triplet-miner tests (CPU and GPU): labels {class, quaternion xyzw, translation} with
duplicated template poses (exact distance ties), a rotation-invariant object, and small random patch pools."""
import numpy as np


def make_labels(seed, n_objects, poses, dup_every=0):
    rng = np.random.default_rng(seed)
    q = rng.standard_normal((n_objects, poses, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=2, keepdims=True).astype(np.float32)
    t = rng.uniform(-0.6, 0.6, (n_objects, poses, 3)).astype(np.float32)
    if dup_every:                                   # pose i + 1 repeats pose i: equal distances to every query
        q[:, 1::dup_every] = q[:, 0:-1:dup_every][:, : q[:, 1::dup_every].shape[1]]
        t[:, 1::dup_every] = t[:, 0:-1:dup_every][:, : t[:, 1::dup_every].shape[1]]
    lab = np.zeros((n_objects, poses, 8), np.float32)
    lab[..., 0] = np.arange(n_objects, dtype=np.float32)[:, None]
    lab[..., 1:5] = q
    lab[..., 5:8] = t
    return lab.reshape(-1, 8)


def make_pool(seed, n, channels, zero_depth_frac=0.0):
    rng = np.random.default_rng(seed)
    p = rng.uniform(0.05, 1.0, (n, 64, 64, channels)).astype(np.float32)
    if zero_depth_frac and channels > 3:
        hole = rng.uniform(size=(n, 64, 64)) < zero_depth_frac
        p[..., 3][hole] = 0.0
    return p


def make_knn_table(seed, max_sim, n_objects, n_tmpl, n_train, hit_frac=0.5):
    """a maxSimKNNTmpl table in pe_knn_bootstrap's layout: some queries hit their puller, some miss with / without a
    second (other-object) entry"""
    rng = np.random.default_rng(seed)
    nq = n_objects * n_train
    table = np.full((nq, 4), -1, np.int32)
    length = np.full(nq, 2, np.int32)
    for q in range(nq):
        obj = q // n_train
        if rng.uniform() < hit_frac:
            table[q, 0], table[q, 1] = obj, max_sim[q, 0]
        else:
            table[q, 0], table[q, 1] = rng.integers(n_objects), rng.integers(n_tmpl)
        if rng.uniform() < 0.5:
            other = (obj + 1 + rng.integers(n_objects - 1)) % n_objects
            table[q, 2], table[q, 3] = other, rng.integers(n_tmpl)
            length[q] = 4
    return table, length
