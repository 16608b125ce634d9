"""Triplet-miner oracle (oracle/oracle_miner.c) against independent numpy restatements and the constraints the
reference's buildBatch guarantees (src/networksolver.cpp:52-186, src/datasetmanager.cpp:687-783).  The reference draws
from std::random_device, so its random picks are pinned by their support and distribution, not by value.  CPU only."""
import numpy as np

from tests import miner_cases as mc


def np_max_sim(tl, ql, n_objects, inplane):
    n_t, n_q = tl.shape[0] // n_objects, ql.shape[0] // n_objects
    out = np.full((n_objects * n_q, 2), -1, np.int32)
    f32 = np.float32
    for o in range(n_objects):
        T = tl[o * n_t:(o + 1) * n_t]
        for p in range(n_q):
            a = ql[o * n_q + p]
            if inplane:
                d = a[1] * T[:, 1]
                for j in (2, 3, 4):
                    d = (d + a[j] * T[:, j]).astype(f32)
                d = np.minimum(np.abs(d), f32(1))
                key = (f32(2) * np.arccos(d.astype(np.float64)).astype(f32)).astype(f32)
                ok = np.ones(n_t, bool)
            else:
                d = a[5] * T[:, 5]
                d = (d + a[6] * T[:, 6]).astype(f32)
                d = (d + a[7] * T[:, 7]).astype(f32)
                key = -d
                ok = d > np.finfo(f32).tiny
            if not ok.any():
                continue
            cand = np.where(ok)[0]
            b = cand[np.flatnonzero(key[cand] == key[cand].min())[0]]          # lowest index among equal keys
            out[o * n_q + p, 0] = b
            cand2 = cand[key[cand] != key[b]]
            if cand2.size:
                out[o * n_q + p, 1] = cand2[np.flatnonzero(key[cand2] == key[cand2].min())[0]]
    return out


def test_max_sim_matches_numpy_including_ties(po):
    for inplane in (True, False):
        tl = mc.make_labels(11, 4, 97, dup_every=5)
        ql = mc.make_labels(12, 4, 33)
        ql[5, 1:5] = tl[97 + 10, 1:5]          # a query equal to a (duplicated) template pose: distance 0 twice
        got = po.max_sim_tmpl(tl, ql, 4, inplane)
        want = np_max_sim(tl, ql, 4, inplane)
        np.testing.assert_array_equal(got, want)
        assert (got[:, 0] >= 0).all() or not inplane
    # translations that never give a positive dot product: the reference leaves both entries unset
    tl = mc.make_labels(13, 2, 9)
    ql = mc.make_labels(14, 2, 3)
    tl[:, 5:8] = np.abs(tl[:, 5:8])
    ql[:, 5:8] = -np.abs(ql[:, 5:8])
    assert (po.max_sim_tmpl(tl, ql, 2, False) == -1).all()


def elevations(po, tl):
    return np.array([po.elevation(l) for l in tl], np.float32)


def test_triplet_constraints_without_bootstrapping(po):
    n_obj, n_t, n_q = 5, 120, 40
    tl, ql = mc.make_labels(21, n_obj, n_t), mc.make_labels(22, n_obj, n_q)
    ms = po.max_sim_tmpl(tl, ql, n_obj, True)
    rot_inv = np.array([0, 1, 0, 2, 0], np.int32)
    el = elevations(po, tl)
    batch = 100                                      # 20 triplets per call, 10 calls = the whole training set
    seen = []
    for it in range(n_obj * n_q // (batch // 5)):
        idx, flag = po.triplet_indices(tl, ms, None, None, rot_inv, None, n_obj, n_q, batch, 5, it, seed=77)
        assert not flag.any()
        for i, row in enumerate(idx):
            lin = it * batch // 5 + i
            pose, obj = lin // n_obj, lin % n_obj
            assert row[0] == obj * n_q + pose                              # anchor (:66-70)
            assert row[1] == obj * n_t + ms[row[0], 0]                     # puller (:73)
            assert row[2] // n_t == obj and row[3] // n_t == obj           # pusher0 / 1: same object
            assert row[4] // n_t != obj                                    # pusher2: another object (:163-165)
            if rot_inv[obj]:
                assert abs(float(el[row[2]]) - float(el[row[1]])) >= 0.2 and abs(float(el[row[3]]) - float(el[row[1]])) >= 0.2
            else:
                assert row[2] != row[1]
            seen.append(row)
    seen = np.array(seen)
    # support and rough uniformity of the random picks: pusher2's object over the 4 admissible ones
    other = (seen[:, 4] // n_t - seen[:, 0] // n_q) % n_obj
    cnt = np.bincount(other, minlength=n_obj)
    assert cnt[0] == 0 and cnt[1:].min() > 0.6 * cnt[1:].mean()
    # the same call is reproducible, another seed is not the same batch
    a, _ = po.triplet_indices(tl, ms, None, None, rot_inv, None, n_obj, n_q, batch, 5, 3, seed=77)
    b, _ = po.triplet_indices(tl, ms, None, None, rot_inv, None, n_obj, n_q, batch, 5, 3, seed=77)
    c, _ = po.triplet_indices(tl, ms, None, None, rot_inv, None, n_obj, n_q, batch, 5, 3, seed=78)
    np.testing.assert_array_equal(a, b)
    assert (a[:, 2:] != c[:, 2:]).any() and (a[:, :2] == c[:, :2]).all()


def test_triplet_bootstrapping_and_hard_negatives(po):
    n_obj, n_t, n_q = 4, 60, 25
    tl, ql = mc.make_labels(31, n_obj, n_t), mc.make_labels(32, n_obj, n_q)
    ms = po.max_sim_tmpl(tl, ql, n_obj, True)
    table, length = mc.make_knn_table(33, ms, n_obj, n_t, n_q)
    rot_inv = np.zeros(n_obj, np.int32)
    neg_off = np.array([0, 0, 7, 7, 12], np.int64)           # objects 1 and 3 have hard negatives
    idx, flag = po.triplet_indices(tl, ms, table, length, rot_inv, neg_off, n_obj, n_q, n_obj * n_q * 5, 5, 0, seed=5)
    assert not flag.any() and idx.shape == (n_obj * n_q, 5)
    n_missed = 0
    for row in idx:
        q = row[0]
        obj = q // n_q
        missed = table[q, 0] != obj or table[q, 1] != ms[q, 0]
        n_missed += missed
        if missed:
            assert row[3] == table[q, 0] * n_t + table[q, 1]                       # :109-111
            if length[q] > 2:
                assert row[4] == table[q, 2] * n_t + table[q, 3]                   # :132-136
            else:
                assert row[4] // n_t != obj
        else:
            assert row[3] // n_t == obj and row[4] // n_t != obj
        if neg_off[obj + 1] > neg_off[obj]:
            assert row[2] < 0 and neg_off[obj] <= ~row[2] < neg_off[obj + 1]       # :95-100
        else:
            assert row[2] >= 0 and row[2] // n_t == obj and row[2] != row[1]
    assert 0 < n_missed < n_obj * n_q


def test_rejection_loops_that_never_end_are_flagged(po):
    # every template of the rotation-invariant object sits at the same elevation
    tl, ql = mc.make_labels(41, 2, 8), mc.make_labels(42, 2, 4)
    tl[:8, 1:5] = tl[0, 1:5]
    ms = po.max_sim_tmpl(tl, ql, 2, True)
    idx, flag = po.triplet_indices(tl, ms, None, None, np.array([1, 0], np.int32), None, 2, 4, 40, 5, 0, seed=1)
    assert (flag[0::2] == 1).all() and (flag[1::2] == 0).all()


def test_gather_is_the_caffe_layout_and_colour_fill_mask(po):
    n_obj, n_t, n_q, ch = 3, 6, 4, 7
    tp, qp = mc.make_pool(51, n_obj * n_t, ch), mc.make_pool(52, n_obj * n_q, ch, zero_depth_frac=0.02)
    neg = mc.make_pool(53, 5, ch)
    idx = np.array([[0, 3, ~2, 7, 12], [5, 1, 4, ~0, 17], [11, 17, 16, 15, 0]], np.int32)
    idx[1, 3] = 9                                           # only pusher0 may be a hard negative in the reference
    out = po.gather_batch(tp, qp, neg, idx, first_linear_id=0, seed=9)
    for t in range(3):
        for s in range(5):
            i = idx[t, s]
            src = qp[i] if s == 0 else (neg[~i] if i < 0 else tp[i])
            np.testing.assert_array_equal(out[t * 5 + s], src.transpose(2, 0, 1))
    filled = po.gather_batch(tp, qp, neg, idx, first_linear_id=40, seed=9, random_background=1)
    for t in range(3):
        zero = qp[idx[t, 0]][..., 3] == 0
        pad = np.pad(zero, 1)
        mask = np.zeros_like(zero)
        for dy in range(3):
            for dx in range(3):
                mask |= pad[dy:dy + 64, dx:dx + 64]
        a, b = filled[t * 5], out[t * 5]
        assert mask.any() and not mask.all()
        np.testing.assert_array_equal(a[:, ~mask], b[:, ~mask])
        v = a[:, mask]
        assert (v >= 0).all() and (v < 1).all() and abs(float(v.mean()) - 0.5) < 0.02 and len(np.unique(v)) > 0.99 * v.size
        np.testing.assert_array_equal(filled[t * 5 + 1:t * 5 + 5], out[t * 5 + 1:t * 5 + 5])
