"""The C++ mirror of the reference's sample producers (include/datasetmanager_b200.hpp: sz::Model::loadModel,
sz::SphereRenderer::createViews, sz::datasetManager::createSynthSamplesWadim / Paul, extractRealSamplesPaul / Wadim)
driven through libpose_b200_host.so the way datasetManager::generateAndStoreSamples drives the originals
(src/datasetmanager.cpp:559-586): model file in, vector<Sample> out.  Checked against the C-ABI chain, which
tests/test_gpu_render.py and tests/test_gpu_patch.py hold bit-exact against the oracle."""
import ctypes as C

import numpy as np
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu
CAM9 = np.array([572.4114, 0, 325.2611, 0, 573.57043, 242.04899, 0, 0, 1], np.float32)
CAM4 = np.array([CAM9[0], CAM9[4], CAM9[2], CAM9[5]], np.float32)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def model(tmp_path_factory):
    """a bumpy coloured sphere with texture coordinates, off-centre so that Model::centroid matters"""
    v, f = scenes.icosphere(3, 0.08)
    v = (v * (1 + 0.12 * np.sin(6 * v[:, 1:2])) + np.float32([0.01, -0.015, 0.02])).astype(np.float32)
    col = ((v / 0.25 + 0.5).clip(0, 1) * 255).astype(np.uint8)
    uv = np.stack([np.arctan2(v[:, 1], v[:, 0]) / np.pi, v[:, 2] / 0.08], 1).astype(np.float32)
    path = tmp_path_factory.mktemp("model") / "obj.ply"
    scenes.write_ply(path, v, f, col, uv, binary=True)
    return str(path).encode()


def _synth(gpu, model, cap, **kw):
    H = gpu.host_lib()
    patches, labels = np.empty((cap, 64, 64, 7), np.float32), np.empty((cap, 8), np.float32)
    n = C.c_int64()
    tex = kw.get("tex")
    gpu.host_check(H.pe_host_create_synth_samples(model, _p(CAM9), kw.get("index", 4), kw.get("subdiv", 0), int(kw.get("inplane", False)),
                                                  int(kw.get("paul", False)), kw.get("dataset_path"), kw.get("dataset_name"), _p(tex),
                                                  tex.shape[0] if tex is not None else 0, tex.shape[1] if tex is not None else 0,
                                                  _p(patches), _p(labels), cap, C.byref(n)))
    return patches[:n.value], labels[:n.value]


def test_create_synth_samples_wadim(gpu, model):
    m = gpu.load_ply(model.decode())
    for inplane, rots in ((False, (0, 15, 10)), (True, (-45, 15, 45))):
        poses = gpu.sphere_views(0, (0.4, 1.1, 1.0), rots)
        assert len(poses) == (7 if inplane else 1) * len(gpu.sphere_views(0, (0.4, 1.1, 1.0), (0, 15, 10)))
        want_p, want_l, status = gpu.template_samples(m["vertices"], m["colors"], m["faces"], CAM4, poses, 4)
        assert not status.any()
        got_p, got_l = _synth(gpu, model, len(poses) + 3, inplane=inplane)
        assert got_p.shape == want_p.shape and np.array_equal(got_p, want_p) and np.array_equal(got_l, want_l)
        assert got_p[..., 3].any() and got_p[..., :3].any()              # the object is in every patch
    # textured model: Model::m_tex set -> the GL_MODULATE path
    tex = np.random.default_rng(2).integers(0, 256, (40, 56, 3)).astype(np.uint8)
    poses = gpu.sphere_views(0, (0.4, 1.1, 1.0), (0, 15, 10))
    want_p, want_l, _ = gpu.template_samples(m["vertices"], m["colors"], m["faces"], CAM4, poses, 4, uv=m["uv"], tex=tex)
    got_p, got_l = _synth(gpu, model, len(poses), tex=tex)
    assert np.array_equal(got_p, want_p) and np.array_equal(got_l, want_l)
    assert not np.array_equal(got_p, _synth(gpu, model, len(poses))[0])
    # too small an output array is an error, not an overrun
    with pytest.raises(gpu.PoseError):
        _synth(gpu, model, len(poses) - 1)
    with pytest.raises(gpu.PoseError):
        _synth(gpu, b"/nonexistent/model.ply", 4)


def _paul_poses(elaz, tz, centroid=None):
    """camRot0 * (Rx(el) * Rz(-az)), t = (0, 0, tz); BigBIRD: the object-frame translation shifted by the centroid"""
    out = []
    for el, az in elaz.astype(np.float64):
        rx = lambda a: np.array([[1, 0, 0], [0, np.cos(a), -np.sin(a)], [0, np.sin(a), np.cos(a)]])
        rz = lambda a: np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
        r = rx(np.pi / 2) @ (rx(el) @ rz(-az))
        t = np.array([0, 0, tz])
        if centroid is not None:                  # pose^-1 = [R^T | -R^T t]; += centroid; invert back
            t = -r @ (-r.T @ t + centroid)
        out.append(np.concatenate([r, t[:, None]], 1))
    return np.asarray(out, np.float32)


def test_create_synth_samples_paul(gpu, model, tmp_path):
    m = gpu.load_ply(model.decode())
    elaz = np.array([[0.3, 0.1], [0.9, 2.0], [1.4, -1.0], [0.05, 4.0]], np.float32)
    (tmp_path / "paul").mkdir()
    np.savetxt(tmp_path / "paul" / "camPositionsElAz.txt", elaz, fmt="%.7f")
    root = (str(tmp_path) + "/").encode()
    for name, tz, cen in ((b"LineMOD", 0.4, None), (b"BigBIRD", 0.72, m["centroid"].astype(np.float64))):
        got_p, got_l = _synth(gpu, model, 8, paul=True, dataset_path=root, dataset_name=name)
        poses = _paul_poses(elaz, tz, cen)
        want_p, want_l, _ = gpu.template_samples(m["vertices"], m["colors"], m["faces"], CAM4, poses, 4,
                                                 z_override=np.full(len(poses), tz, np.float32))
        assert got_p.shape == want_p.shape
        # the poses are built in fp32 by the mirror and in fp64 here: labels agree to rounding, patches up to the few pixels
        # whose coverage flips with a 1-ulp pose change
        np.testing.assert_allclose(got_l, want_l, rtol=0, atol=2e-6)
        assert np.mean(np.abs(got_p - want_p) > 1e-3) < 0.01
        assert got_p[..., 3].any()
    with pytest.raises(gpu.PoseError):            # camPositionsElAz.txt missing
        _synth(gpu, model, 8, paul=True, dataset_path=b"/nonexistent/")


def test_extract_real_samples(gpu, model):
    m = gpu.load_ply(model.decode())
    H = gpu.host_lib()
    # "real" frames: renders of the model at poses off the image centre, in front of nothing
    poses = gpu.sphere_views(0, (0.5, 1.1, 1.0), (0, 15, 10))[:5].copy()
    poses[:, 0, 3] += np.float32([0.05, -0.08, 0.0, 0.1, -0.02])
    poses[:, 1, 3] += np.float32([-0.03, 0.02, 0.06, 0.0, -0.05])
    col, dep = gpu.render_views(m["vertices"], m["colors"], m["faces"], CAM4, poses)
    n = len(poses)
    pp = np.ascontiguousarray(poses.reshape(n, 12))

    def run(wadim, name, ply):
        patches, labels = np.empty((n, 64, 64, 7), np.float32), np.empty((n, 8), np.float32)
        cnt = C.c_int64()
        gpu.host_check(H.pe_host_extract_real_samples(_p(col), _p(dep), n, 480, 640, _p(pp), _p(CAM9), 2, int(wadim), ply, name,
                                                      _p(patches), _p(labels), n, C.byref(cnt)))
        assert cnt.value == n
        return patches, labels

    for wadim, name, ply, kw in ((False, b"LineMOD", None, {}), (False, b"BigBIRD", model, {"centroid": m["centroid"]}),
                                 (True, b"LineMOD", None, {"z_from_depth": True})):
        got_p, got_l = run(wadim, name, ply)
        want_p, want_l, status = gpu.scene_samples(col, dep, CAM4, poses, 2, **kw)
        assert not status.any()
        assert np.array_equal(got_p, want_p) and np.array_equal(got_l, want_l)
        assert got_p[..., 3].any()
    # the BigBIRD centroid moves the cut: not the same patches as LineMOD's
    assert not np.array_equal(run(False, b"BigBIRD", model)[0], run(False, b"LineMOD", None)[0])


def test_create_views_keeps_frames(gpu, model):
    m = gpu.load_ply(model.decode())
    H = gpu.host_lib()
    sc, ro = np.float32([0.4, 1.1, 1.0]), np.float32([0, 15, 10])
    cnt = C.c_int64()
    gpu.host_check(H.pe_host_create_views(None, _p(CAM9), 0, _p(sc), _p(ro), 1, None, None, None, 0, C.byref(cnt)))
    poses = gpu.sphere_views(0, sc, ro)
    assert cnt.value == len(poses)
    n = cnt.value
    col, dep, pp = np.empty((n, 480, 640, 3), np.uint8), np.empty((n, 480, 640), np.float32), np.empty((n, 3, 4), np.float32)
    gpu.host_check(H.pe_host_create_views(model, _p(CAM9), 0, _p(sc), _p(ro), 1, _p(col), _p(dep), _p(pp), n, C.byref(cnt)))
    want_c, want_d = gpu.render_views(m["vertices"], m["colors"], m["faces"], CAM4, poses)
    assert np.array_equal(pp, poses) and np.array_equal(col, want_c) and np.array_equal(dep, want_d)
    assert (dep > 0).any(axis=(1, 2)).all()
