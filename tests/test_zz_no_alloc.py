"""pe_histogram_dev_noalloc (no counterpart in the reference): the
computeHistogram reduction on a fixed file-scope device scratch, for callers that cannot call cudaMalloc.

Checked through the C ABI against the existing entry point pe_histogram_dev (and the oracle) on two calls in a row that
differ in size, k, rotInv pattern and neighbours - the shared scratch keeps the first call's contents, so the second
call proves that what must start at zero is cleared - plus the size limits.  Collected last on purpose (file name).
Acceptance as in tests/test_gpu_parity.py: bins exact, mean / median within 1e-3 deg.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BINS = np.array([-1, 0, 10, 20, 40, 180], np.float32)
MAX_QUERIES, MAX_CLASSES = 65536, 1024          # the limits include/pose_b200.h states


def _problem(synth, seed, n_obj, nr_t, nr_q, k, own_frac):
    """random neighbours: a mix of same-class templates and arbitrary ones, all rotInv kinds"""
    tl = synth.make_labels(seed, n_obj, nr_t)
    src = (np.arange(nr_q) * 13) % nr_t
    ql = synth.perturb_labels(tl.reshape(n_obj, nr_t, 8)[:, src].reshape(-1, 8), seed + 1, max_deg=25.0)
    rng = np.random.default_rng(seed)
    nq = n_obj * nr_q
    own = (np.repeat(np.arange(n_obj), nr_q) * nr_t)[:, None] + rng.integers(0, nr_t, (nq, k))
    anyc = rng.integers(0, n_obj * nr_t, (nq, k))
    idx = np.where(rng.random((nq, k)) < own_frac, own, anyc).astype(np.int32)
    return tl, ql, idx, synth.rot_inv_pattern(n_obj)


def test_histogram_noalloc_equals_histogram_dev_twice(gpu, po, synth):
    cases = [_problem(synth, 77, 20, 700, 150, 5, 0.5),      # 3,000 queries, k = 5
             _problem(synth, 91, 7, 60, 11, 3, 0.5)]         # 77 queries, k = 3: fewer queries than the call before it
    for call, (tl, ql, idx, rot_inv) in enumerate(cases):
        n_obj = len(rot_inv)
        nr_t = len(tl) // n_obj
        db = gpu.TemplateDB(desc=np.zeros((4, 16), np.float32), labels_all=tl, poses_per_object=nr_t, t_total=len(tl))
        d_i, d_l = gpu.DeviceArray.from_host(idx), gpu.DeviceArray.from_host(ql)
        nq, k = idx.shape
        want = db.histogram_dev(d_i, d_l, nq, k, rot_inv, BINS)
        got = db.histogram_dev_noalloc(d_i, d_l, nq, k, rot_inv, BINS)
        np.testing.assert_array_equal(got[:6], want[:6], err_msg=f"call {call}")
        np.testing.assert_allclose(got[6:], want[6:], rtol=0, atol=1e-3, err_msg=f"call {call}")
        oracle = po.histogram(idx, ql, tl, nr_t, rot_inv, BINS)
        np.testing.assert_array_equal(got[:6], oracle[:6])
        np.testing.assert_allclose(got[6:], oracle[6:], rtol=0, atol=1e-3)
        assert 0 < got[0] < 1 and got[5] > 0          # a non-degenerate histogram
        # the same call again: identical (nothing accumulates in the shared scratch)
        np.testing.assert_array_equal(db.histogram_dev_noalloc(d_i, d_l, nq, k, rot_inv, BINS), got)
    # nobody matched at all -> NaN median, as the existing entry point
    tl, ql, idx, rot_inv = cases[1]
    q_class = ql[:, 0].astype(np.int64)[:, None]
    other = ((q_class + 1) % 7 * 60 + idx % 60).astype(np.int32)       # every neighbour from the class after the query's
    db = gpu.TemplateDB(desc=np.zeros((4, 16), np.float32), labels_all=tl, poses_per_object=60, t_total=len(tl))
    d_i, d_l = gpu.DeviceArray.from_host(other), gpu.DeviceArray.from_host(ql)
    got = db.histogram_dev_noalloc(d_i, d_l, len(other), 3, rot_inv, BINS)
    np.testing.assert_array_equal(got, db.histogram_dev(d_i, d_l, len(other), 3, rot_inv, BINS))
    assert np.isnan(got[7]) and got[0] == 1.0


def test_histogram_noalloc_rejects_sizes_beyond_its_scratch(gpu, synth):
    tl, ql, idx, rot_inv = _problem(synth, 5, 4, 10, 3, 2, 0.5)
    db = gpu.TemplateDB(desc=np.zeros((4, 16), np.float32), labels_all=tl, poses_per_object=10, t_total=len(tl))
    d_i, d_l = gpu.DeviceArray.from_host(idx), gpu.DeviceArray.from_host(ql)
    with pytest.raises(gpu.PoseError) as e:             # rejected before any pointer is touched
        db.histogram_dev_noalloc(d_i, d_l, MAX_QUERIES + 1, 2, rot_inv, BINS)
    assert e.value.code == gpu.PE_ERR_INVALID and str(MAX_QUERIES) in str(e.value)
    with pytest.raises(gpu.PoseError) as e:
        db.histogram_dev_noalloc(d_i, d_l, len(idx), 2, np.zeros(MAX_CLASSES + 1, np.int32), BINS)
    assert e.value.code == gpu.PE_ERR_INVALID and str(MAX_CLASSES) in str(e.value)
    # exactly at the query limit the call works (the 12 queries tiled up to 65,536)
    reps = MAX_QUERIES // len(idx) + 1
    big_i, big_l = np.tile(idx, (reps, 1))[:MAX_QUERIES], np.tile(ql, (reps, 1))[:MAX_QUERIES]
    d_bi, d_bl = gpu.DeviceArray.from_host(big_i), gpu.DeviceArray.from_host(big_l)
    got = db.histogram_dev_noalloc(d_bi, d_bl, MAX_QUERIES, 2, rot_inv, BINS)
    want = db.histogram_dev(d_bi, d_bl, MAX_QUERIES, 2, rot_inv, BINS)
    np.testing.assert_array_equal(got[:6], want[:6])
    np.testing.assert_allclose(got[6:], want[6:], rtol=0, atol=1e-3)
    # at the limit of classes the call works
    rot_big = np.zeros(MAX_CLASSES, np.int32)
    rot_big[:len(rot_inv)] = rot_inv
    np.testing.assert_array_equal(db.histogram_dev_noalloc(d_i, d_l, len(idx), 2, rot_big, BINS),
                                  db.histogram_dev(d_i, d_l, len(idx), 2, rot_inv, BINS))
