"""GPU parity of the triplet miner through the C ABI: most-similar-template table, triplet selection,
Caffe-layout batch and colour fill bit for bit against oracle/oracle_miner.c, and pe_miner_compute_knn against the same
chain composed from the evaluation entry points (forward -> k = 3 matcher -> bootstrap table)."""
import numpy as np
import pytest

from tests import miner_cases as mc
from tests import netdefs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("inplane", [True, False], ids=["inplane", "translation-dot"])
def test_max_sim_bit_exact(gpu, po, inplane):
    n_obj = 5
    tl = mc.make_labels(11, n_obj, 301, dup_every=7)
    ql = mc.make_labels(12, n_obj, 64)
    ql[9, 1:5] = tl[301 + 14, 1:5]
    m = gpu.Miner(None, tl, None, ql, n_obj, np.zeros(n_obj, np.int32), inplane=inplane, channels=3)
    np.testing.assert_array_equal(m.max_sim(), po.max_sim_tmpl(tl, ql, n_obj, inplane))
    m.close()
    if not inplane:
        tl2, ql2 = tl.copy(), ql.copy()
        tl2[:, 5:8] = np.abs(tl2[:, 5:8])
        ql2[:, 5:8] = -np.abs(ql2[:, 5:8])
        m = gpu.Miner(None, tl2, None, ql2, n_obj, np.zeros(n_obj, np.int32), inplane=False, channels=3)
        assert (m.max_sim() == -1).all()
        with pytest.raises(gpu.PoseError):
            m.build_batch(10, 0, seed=1, want_batch=False)           # no puller: the reference would read garbage
        m.close()


@pytest.mark.parametrize("channels", [3, 4, 7])
def test_build_batch_bit_exact(gpu, po, channels):
    n_obj, n_t, n_q = 4, 40, 15
    tl, ql = mc.make_labels(21, n_obj, n_t), mc.make_labels(22, n_obj, n_q)
    tp, qp = mc.make_pool(23, n_obj * n_t, channels), mc.make_pool(24, n_obj * n_q, channels, zero_depth_frac=0.02)
    rot_inv = np.array([0, 1, 0, 2], np.int32)
    m = gpu.Miner(tp, tl, qp, ql, n_obj, rot_inv, inplane=True)
    ms = po.max_sim_tmpl(tl, ql, n_obj, True)
    batch = 30
    # plain rounds
    for it in (0, 3, n_obj * n_q // 6 - 1):
        got_b, got_i = m.build_batch(batch, it, seed=1234 + it)
        want_i, flag = po.triplet_indices(tl, ms, None, None, rot_inv, None, n_obj, n_q, batch, 5, it, 1234 + it)
        assert not flag.any()
        np.testing.assert_array_equal(got_i, want_i)
        np.testing.assert_array_equal(got_b, po.gather_batch(tp, qp, None, want_i, it * batch // 5, 1234 + it))
    # bootstrapping rounds with hard negatives
    table, length = mc.make_knn_table(25, ms, n_obj, n_t, n_q)
    neg_off = np.array([0, 3, 3, 9, 9], np.int64)
    neg = mc.make_pool(26, 9, channels)
    m.set_knn(table, length)
    m.set_hard_negatives(neg, neg_off)
    fill = 1 if channels == 7 else 0
    got_b, got_i = m.build_batch(n_obj * n_q * 5, 0, seed=99, random_background=fill)
    want_i, flag = po.triplet_indices(tl, ms, table, length, rot_inv, neg_off, n_obj, n_q, n_obj * n_q * 5, 5, 0, 99)
    assert not flag.any() and (want_i[:, 2] < 0).any()
    np.testing.assert_array_equal(got_i, want_i)
    np.testing.assert_array_equal(got_b, po.gather_batch(tp, qp, neg, want_i, 0, 99, fill))
    # switching both off again gives the plain batch back
    m.set_knn(None, None)
    m.set_hard_negatives(None, None)
    _, got_i = m.build_batch(batch, 1, seed=7, want_batch=False)
    want_i, _ = po.triplet_indices(tl, ms, None, None, rot_inv, None, n_obj, n_q, batch, 5, 1, 7)
    np.testing.assert_array_equal(got_i, want_i)
    # argument errors the reference would crash or hang on
    with pytest.raises(gpu.PoseError):
        m.build_batch(batch, n_obj * n_q // 6 + 5, seed=1)           # iter past the training set
    with pytest.raises(gpu.PoseError):
        m.build_batch(batch, 0, seed=1, random_background=3)         # bgfill's image backgrounds are out of scope
    if channels != 7:
        with pytest.raises(gpu.PoseError):
            m.build_batch(batch, 0, seed=1, random_background=1)
    m.close()


def test_endless_rejection_loop_is_an_error(gpu):
    tl, ql = mc.make_labels(41, 2, 8), mc.make_labels(42, 2, 4)
    tl[:8, 1:5] = tl[0, 1:5]
    m = gpu.Miner(None, tl, None, ql, 2, np.array([1, 0], np.int32), channels=3)
    with pytest.raises(gpu.PoseError):
        m.build_batch(40, 0, seed=1, want_batch=False)
    m.close()


def test_compute_knn_equals_the_composed_chain(gpu, po, synth):
    """networkSolver::computeKNN on resident pools == computeDescriptors on both sets + knnMatch(3) + table"""
    name = "manifold_rgbdnor_16"
    ref = po.parse_prototxt(netdefs.prototxt(name))
    net = gpu.Net(text=netdefs.prototxt(name))
    net.set_blobs(synth.make_blobs(ref.blob_names, ref.blob_shapes, 3, mode=1))
    n_obj, n_t, n_q = 3, 50, 20
    tl, ql = mc.make_labels(61, n_obj, n_t), mc.make_labels(62, n_obj, n_q)
    tp = synth.make_patches(5, 0, n_obj * n_t, 7, n_t)
    qp = tp.reshape(n_obj, n_t, 64, 64, 7)[:, :n_q].reshape(-1, 64, 64, 7) + mc.make_pool(64, n_obj * n_q, 7) * np.float32(0.05)
    m = gpu.Miner(tp, tl, qp, ql, n_obj, np.zeros(n_obj, np.int32))
    table, length = m.compute_knn(net)
    dt, dq = net.compute_descriptors(tp), net.compute_descriptors(qp)
    db = gpu.TemplateDB(dt, labels_all=tl, poses_per_object=n_t)
    idx, _ = db.knn(dq, 3)
    want_t, want_l = db.knn_bootstrap(idx, n_q)
    np.testing.assert_array_equal(table, want_t)
    np.testing.assert_array_equal(length, want_l)
    o_t, o_l = po.knn_bootstrap(idx, n_q, n_t)
    np.testing.assert_array_equal(table, o_t)
    np.testing.assert_array_equal(length, o_l)
    # the table now drives the selection (bootstrapping switched on by compute_knn)
    ms = po.max_sim_tmpl(tl, ql, n_obj, True)
    _, got_i = m.build_batch(n_obj * n_q * 5, 0, seed=3, want_batch=False)
    want_i, _ = po.triplet_indices(tl, ms, table, length, np.zeros(n_obj, np.int32), None, n_obj, n_q, n_obj * n_q * 5, 5, 0, 3)
    np.testing.assert_array_equal(got_i, want_i)
    m.close()
