"""The C++ host mirror (include/networkevaluator_b200.hpp: sz::networkEvaluator, sz::networkSolver) driven through the
flat-array doors of libpose_b200_host.so the way a C++ caller of the reference drives the originals
(vector<vector<Sample>> sets, static methods, the INI config), against the C-ABI chain and the CPU oracle."""
import ctypes as C
import os

import numpy as np
import pytest

from tests import miner_cases as mc
from tests import netdefs

pytestmark = pytest.mark.gpu
NAME = "manifold_rgbdnor_16"
N_OBJ, N_T, N_Q = 3, 40, 12
BINS = np.array([-1, 0, 10, 20, 40, 180], np.float32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def world(gpu, po, synth):
    ref = po.parse_prototxt(netdefs.prototxt(NAME))
    net = gpu.Net(text=netdefs.prototxt(NAME))
    net.set_blobs(synth.make_blobs(ref.blob_names, ref.blob_shapes, 3, mode=1))
    tl = synth.make_labels(5, N_OBJ, N_T)
    tp = synth.make_patches(5, 0, N_OBJ * N_T, 7, N_T)
    qp = (tp.reshape(N_OBJ, N_T, 64, 64, 7)[:, 3:3 + N_Q].reshape(-1, 64, 64, 7) + mc.make_pool(64, N_OBJ * N_Q, 7) * np.float32(0.05))
    ql = synth.perturb_labels(tl.reshape(N_OBJ, N_T, 8)[:, 3:3 + N_Q].reshape(-1, 8), 9)
    return dict(net=net, tp=np.ascontiguousarray(tp), tl=np.ascontiguousarray(tl, np.float32), qp=np.ascontiguousarray(qp),
                ql=np.ascontiguousarray(ql, np.float32))


def test_compute_descriptors_and_histogram_and_confusion(gpu, po, world):
    H = gpu.host_lib()
    net, tp, tl, qp, ql = (world[k] for k in ("net", "tp", "tl", "qp", "ql"))
    out = np.empty((len(qp), 16), np.float32)
    gpu.host_check(H.pe_host_compute_descriptors(net.h, _p(qp), _p(ql), len(qp), 7, _p(out)))
    dq = net.compute_descriptors(qp)
    np.testing.assert_array_equal(out, dq)
    dt = net.compute_descriptors(tp)
    db = gpu.TemplateDB(dt, labels_all=tl, poses_per_object=N_T)
    rot_inv = np.array([0, 1, 2], np.int32)
    for knn in (1, 3):
        idx, _ = db.knn(dq, knn)
        hist = np.empty(len(BINS) + 2, np.float32)
        gpu.host_check(H.pe_host_compute_histogram(net.h, _p(tp), _p(tl), N_OBJ, N_T, _p(qp), _p(ql), N_Q, 7, _p(rot_inv), 3,
                                                   _p(BINS), len(BINS), knn, _p(hist)))
        np.testing.assert_array_equal(hist, db.histogram(idx, ql, rot_inv, BINS))
        np.testing.assert_allclose(hist, po.histogram(idx, ql, tl, N_T, rot_inv, BINS), rtol=0, atol=1e-6)
        cm = np.empty((3, 3), np.float32)
        gpu.host_check(H.pe_host_compute_confusion(net.h, _p(tp), _p(tl), N_OBJ, N_T, _p(qp), _p(ql), N_Q, 7, b"ape, bowl, cam, duck",
                                                   b"cam,ape,bowl", knn, _p(cm)))
        np.testing.assert_array_equal(cm, db.confusion(idx, ql, np.array([1, 2, 0, -1], np.int32), 3, N_Q))
        np.testing.assert_array_equal(cm, po.confusion(idx, ql, tl, np.array([1, 2, 0, -1], np.int32), 3, N_Q))
        assert abs(float(cm.sum()) - 3.0) < 1e-5


def test_solver_compute_knn_and_accuracy(gpu, po, world):
    H = gpu.host_lib()
    net, tp, tl, qp, ql = (world[k] for k in ("net", "tp", "tl", "qp", "ql"))
    ms = po.max_sim_tmpl(tl, ql, N_OBJ, True)
    ms0 = np.ascontiguousarray(ms[:, 0])
    table, length = np.empty((len(qp), 4), np.int32), np.empty(len(qp), np.int32)
    intra, inter = C.c_float(), C.c_float()
    gpu.host_check(H.pe_host_compute_knn(net.h, _p(tp), _p(tl), N_OBJ, N_T, _p(qp), _p(ql), N_Q, 7, _p(ms0), _p(table), _p(length),
                                         C.byref(intra), C.byref(inter)))
    db = gpu.TemplateDB(net.compute_descriptors(tp), labels_all=tl, poses_per_object=N_T)
    idx, _ = db.knn(net.compute_descriptors(qp), 3)
    want_t, want_l = po.knn_bootstrap(idx, N_Q, N_T)
    np.testing.assert_array_equal(table, want_t)
    np.testing.assert_array_equal(length, want_l)
    o_intra, o_inter = po.knn_accuracy(ms0, want_t, N_Q)
    assert intra.value == pytest.approx(o_intra, abs=1e-4) and inter.value == pytest.approx(o_inter, abs=1e-4)
    assert inter.value > 90          # the queries are noisy copies of templates of their own object


def test_solver_build_batch(gpu, po, world):
    H = gpu.host_lib()
    tp, tl, qp, ql = (world[k] for k in ("tp", "tl", "qp", "ql"))
    rot_inv = np.array([0, 1, 0], np.int32)
    ms = po.max_sim_tmpl(tl, ql, N_OBJ, True)
    table, length = mc.make_knn_table(7, ms, N_OBJ, N_T, N_Q)
    batch = 30
    for boot in (False, True):
        out_b = np.empty((batch, 7, 64, 64), np.float32)
        out_s = np.empty((batch, 64, 64, 7), np.float32)
        out_i, out_ms = np.empty((batch // 5, 5), np.int32), np.empty((N_OBJ * N_Q, 2), np.int32)
        gpu.host_check(H.pe_host_build_batch(_p(tp), _p(tl), N_OBJ, N_T, _p(qp), _p(ql), N_Q, 7, _p(rot_inv), 1,
                                             _p(table) if boot else None, _p(length) if boot else None, batch, 2, 4242, 0,
                                             _p(out_b), _p(out_i), _p(out_ms), _p(out_s)))
        np.testing.assert_array_equal(out_ms, ms)
        want_i, flag = po.triplet_indices(tl, ms, table if boot else None, length if boot else None, rot_inv, None, N_OBJ, N_Q,
                                          batch, 5, 2, 4242)          # first batch of a solver: seed + 0
        assert not flag.any()
        np.testing.assert_array_equal(out_i, want_i)
        np.testing.assert_array_equal(out_b, po.gather_batch(tp, qp, None, want_i, 2 * batch // 5, 4242))
        np.testing.assert_array_equal(out_s.transpose(0, 3, 1, 2), out_b)        # vector<Sample> form == Caffe form


def test_save_logs_writes_the_reference_files(gpu, world, tmp_path):
    H = gpu.host_lib()
    net, tp, tl, qp, ql = (world[k] for k in ("net", "tp", "tl", "qp", "ql"))
    cfg = tmp_path / "config.ini"
    cfg.write_text(f"""[paths]
output_path = {tmp_path}/

[input]
models = ape,bowl,cam
used_models = ape,bowl,cam
rotInv = 0,1,2
random_background = 0

[output]
kNN = 1
log_name = _t.log
""")
    for it in (60, 120):
        gpu.host_check(H.pe_host_save_logs(net.h, _p(tp), _p(tl), N_OBJ, N_T, _p(qp), _p(ql), N_Q, _p(qp), _p(ql), N_Q, 7,
                                           os.fsencode(str(cfg)), it, 1.5))
    log = (tmp_path / "log_t.log").read_text().splitlines()
    assert log[0].split("\t")[:4] == ["mod", "iter", "epoch", "time"] and len(log) == 5
    assert [l.split("\t")[0] for l in log[1:]] == ["test", "train", "test", "train"]
    cm = (tmp_path / "cm_t.log").read_text().splitlines()
    assert cm[0].split("\t") == ["ape", "bowl", "cam"] and len(cm) == 4
    diag = [float(cm[1 + i].split("\t")[1 + i]) for i in range(3)]
    assert min(diag) > 90
