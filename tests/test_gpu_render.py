"""GPU parity of the template renderer through the C ABI: pe_sphere_views / pe_render_views against
the CPU oracle (bit-exact: integer coverage, identical float op order), then views -> patches -> descriptors on device."""
import numpy as np
import pytest

from tests import netdefs, scenes

pytestmark = pytest.mark.gpu
CAM = np.array([572.4114, 573.57043, 325.2611, 242.04899], np.float32)


def test_sphere_views_equal_oracle(gpu, po):
    for depth, scale, rot, skip in ((2, (0.4, 0.2, 0.8), (-45, 15, 45), True), (1, (0.5, 1.0, 0.5), (0, 30, 330), False), (0, (1, 1, 1), (0, 1, 0), True)):
        got = gpu.sphere_views(depth, scale, rot, skip)
        want = po.sphere_views(depth, scale, rot, skip)
        assert got.shape == want.shape and np.array_equal(got, want)


def test_render_bit_exact_vs_oracle(gpu, po):
    rng = np.random.default_rng(3)
    v, f = scenes.icosphere(3, 0.09)
    v = (v * (1 + 0.15 * np.sin(7 * v[:, :1]))).astype(np.float32)            # a bumpy blob: self-occlusion near the silhouette
    c = rng.random(v.shape).astype(np.float32)
    cv, cc, cf = scenes.cube(0.05)
    meshes = [(v, c, f), (cv + np.float32([0.02, -0.01, 0.0]), cc, cf)]
    poses = po.sphere_views(1, (0.45, 0.35, 0.8), (-45, 45, 45))[::9]
    for vv, col, ff in meshes:
        gc, gd = gpu.render_views(vv, col, ff, CAM, poses)
        assert gc.shape == (len(poses), 480, 640, 3)
        for i, rt in enumerate(poses):
            oc, od = po.render(vv, col, ff, CAM, rt)
            assert np.array_equal(gd[i], od), i
            assert np.array_equal(gc[i], oc), i
            assert (od > 0).sum() > 3000
    # degenerate inputs: no faces, faces with bad indices, object behind the camera
    gc, gd = gpu.render_views(v, c, np.zeros((0, 3), np.int32), CAM, poses[:1])
    assert not gd.any() and not gc.any()
    bad = np.array([[0, 1, 99999], [-1, 2, 3]], np.int32)
    assert not gpu.render_views(v, c, bad, CAM, poses[:1])[1].any()
    behind = poses[:1].copy(); behind[0, 2, 3] = -1.0
    assert not gpu.render_views(v, c, f, CAM, behind)[1].any()


def test_views_to_descriptors_on_device(gpu, po, synth):
    """the template pass of datasetManager::createTemplatesPaul (src/datasetmanager.cpp:505-523) on the device:
    render -> patch at the principal point, z = pose distance -> descriptor; equal to the chain through host buffers"""
    v, f = scenes.icosphere(3, 0.08)
    c = (v / 0.16 + 0.5).astype(np.float32)
    poses = gpu.sphere_views(1, (0.5, 0.2, 0.7), (-30, 30, 30))
    n = len(poses)
    ref = po.parse_prototxt(netdefs.prototxt("manifold_rgbdnor_16"))
    blobs = synth.make_blobs(ref.blob_names, ref.blob_shapes, 4, mode=1)
    net = gpu.Net(text=netdefs.prototxt("manifold_rgbdnor_16"))
    net.set_blobs(blobs)
    d_v, d_c, d_f = gpu.DeviceArray.from_host(v), gpu.DeviceArray.from_host(c), gpu.DeviceArray.from_host(f)
    d_pose = gpu.DeviceArray.from_host(poses)
    d_col = gpu.DeviceArray((n, 480, 640, 3), np.uint8)
    d_dep = gpu.DeviceArray((n, 480, 640))
    ctr = np.tile(np.array([[int(CAM[2]), int(CAM[3])]], np.int32), (n, 1))    # cam(0,2), cam(1,2) truncated to int as the call does
    z = poses[:, 2, 3].copy()
    d_ctr, d_z = gpu.DeviceArray.from_host(ctr), gpu.DeviceArray.from_host(z)
    d_p = gpu.DeviceArray((n, 64, 64, 7))
    d_desc = gpu.DeviceArray((n, 16))
    gpu.render_views_dev(net, d_v, d_c, len(v), d_f, len(f), CAM, d_pose, n, 480, 640, d_col, d_dep)
    gpu.sample_patches_dev(net, d_col, d_dep, None, n, 480, 640, None, d_ctr, d_z, n, float(CAM[0]), float(CAM[1]), 0.2, d_p)
    net.compute_descriptors_dev(d_p, n, d_desc)
    net.sync()
    col, dep = gpu.render_views(v, c, f, CAM, poses)
    assert np.array_equal(d_dep.to_host(), dep) and np.array_equal(d_col.to_host(), col)
    patches, status = gpu.sample_patches(col, dep, ctr, z, float(CAM[0]), float(CAM[1]))
    assert not status.any() and np.array_equal(d_p.to_host(), patches)
    assert np.array_equal(d_desc.to_host(), net.compute_descriptors(patches))
    # the object sits in the middle of the patch: its depth channel is 0.5 +- the radius / (2 m), colour is non-black there
    mid = patches[:, 24:40, 24:40]
    assert np.all(np.abs(mid[..., 3] - 0.5) < 0.3) and np.all(mid[..., :3].max(-1) > 0)
    # distinct views give distinct descriptors
    dd = d_desc.to_host()
    assert len(np.unique(np.round(dd, 4), axis=0)) == n


def test_near_plane_clipping_bit_exact_vs_oracle(gpu, po):
    """triangles through the near plane are clipped, not dropped: the analytic ground-plane scene of
    tests/test_render_oracle.py and a mesh the camera sits inside, CUDA == oracle bit for bit"""
    v, c, f = scenes.ground_plane()
    rt = np.zeros((1, 3, 4), np.float32)
    rt[0, :, :3] = np.eye(3)
    gc, gd = gpu.render_views(v, c, f, CAM, rt, z_near=0.3, z_far=4.0)
    oc, od = po.render(v, c, f, CAM, rt[0], z_near=0.3, z_far=4.0)
    assert np.array_equal(gd[0], od) and np.array_equal(gc[0], oc) and (od > 0).sum() > 100000
    # camera inside a bumpy sphere: most triangles straddle or lie behind the near plane somewhere
    sv, sf = scenes.icosphere(3, 0.5)
    sc = (sv + 0.5).astype(np.float32)
    poses = po.sphere_views(0, (0.05, 0.2, 0.45), (0, 90, 90))[::5]
    g2c, g2d = gpu.render_views(sv, sc, sf, CAM, poses, z_near=0.2, z_far=4.0)
    for i, p in enumerate(poses):
        oc, od = po.render(sv, sc, sf, CAM, p, z_near=0.2, z_far=4.0)
        assert np.array_equal(g2d[i], od) and np.array_equal(g2c[i], oc), i
    assert (g2d > 0).mean() > 0.5


def test_textured_mesh_bit_exact_vs_oracle(gpu, po, tmp_path):
    """per-vertex uv + BGR texture, GL_MODULATE / GL_REPEAT / GL_LINEAR (reference src/model.cpp:330-350,
    src/painter.cpp:140-160): the analytic quad and a textured blob loaded through the PLY reader"""
    v, c, uv, f, tex = scenes.textured_quad()
    rt = np.zeros((1, 3, 4), np.float32)
    rt[0, :, :3] = np.eye(3)
    gc, gd = gpu.render_views_tex(v, c, uv, f, tex, CAM, rt)
    oc, od = po.render(v, c, f, CAM, rt[0], uv=uv, tex=tex)
    assert np.array_equal(gd[0], od) and np.array_equal(gc[0], oc)
    rng = np.random.default_rng(9)
    bv, bf = scenes.icosphere(3, 0.09)
    buv = (np.stack([np.arctan2(bv[:, 1], bv[:, 0]) / np.pi, bv[:, 2] / 0.09], 1) * 1.5).astype(np.float32)
    scenes.write_ply(tmp_path / "blob.ply", bv, bf, rng.integers(60, 256, (len(bv), 3)), buv, binary=True)
    m = gpu.load_ply(tmp_path / "blob.ply")
    big = rng.integers(0, 256, (64, 96, 3)).astype(np.uint8)
    poses = po.sphere_views(1, (0.4, 0.3, 0.7), (-30, 60, 30))[::11]
    gc, gd = gpu.render_views_tex(m["vertices"], m["colors"], m["uv"], m["faces"], big, CAM, poses)
    for i, p in enumerate(poses):
        oc, od = po.render(m["vertices"], m["colors"], m["faces"], CAM, p, uv=m["uv"], tex=big)
        assert np.array_equal(gd[i], od) and np.array_equal(gc[i], oc), i
    plain = gpu.render_views(m["vertices"], m["colors"], m["faces"], CAM, poses[:1])[0]
    assert not np.array_equal(plain[0], gc[0])
    with pytest.raises(gpu.PoseError):      # uv without a texture
        gpu.lib().pe_render_views_tex  # symbol exists
        gpu.check(gpu.lib().pe_render_views_tex(0, m["vertices"].ctypes.data, m["colors"].ctypes.data, m["uv"].ctypes.data, len(m["vertices"]),
                                                m["faces"].ctypes.data, len(m["faces"]), None, 0, 0, CAM.ctypes.data, poses.ctypes.data, 1, 480, 640,
                                                0.001, 4.0, plain.ctypes.data, gd.ctypes.data))


def test_template_samples_equal_the_oracle_chain(gpu, po, synth):
    """pe_template_samples = datasetManager::createSynthSamplesWadim (src/datasetmanager.cpp:498-535): view -> render ->
    normals -> patch at the principal point with z = pose t_z -> label; bit-exact against the oracle's stages chained by
    hand, for an untextured and a textured mesh, and the device variant feeds the forward without a host round trip"""
    rng = np.random.default_rng(6)
    v, f = scenes.icosphere(3, 0.08)
    v = (v * (1 + 0.12 * np.sin(6 * v[:, 1:2]))).astype(np.float32)
    c = (v / 0.2 + 0.5).clip(0, 1).astype(np.float32)
    uv = (np.stack([np.arctan2(v[:, 1], v[:, 0]) / np.pi, v[:, 2] / 0.08], 1)).astype(np.float32)
    tex = rng.integers(0, 256, (48, 64, 3)).astype(np.uint8)
    poses = gpu.sphere_views(1, (0.4, 0.35, 1.1), (-45, 45, 45))[::7]       # the reference's scales (0.4, 1.1, 1.0) -> one distance
    n = len(poses)
    cx, cy = int(CAM[2]), int(CAM[3])
    for kw in ({}, {"uv": uv, "tex": tex}):
        patches, labels, status = gpu.template_samples(v, c, f, CAM, poses, 3, **kw)
        assert patches.shape == (n, 64, 64, 7) and not status.any()
        for i, rt in enumerate(poses):
            col, dep = po.render(v, c, f, CAM, rt, **kw)
            nor = po.depth2normals(dep, float(CAM[0]), float(CAM[1]))
            want = po.sample_patch(col, dep, nor, cx, cy, float(rt[2, 3]), float(CAM[0]), float(CAM[1]))
            assert np.array_equal(patches[i], want), i
            assert np.array_equal(labels[i], po.pose_label(rt, 3))
        assert np.all(labels[:, 0] == 3) and np.abs(np.linalg.norm(labels[:, 1:5], axis=1) - 1).max() < 1e-6
    # z override (the Paul variant passes trans(2)): a nearer z cuts a larger rectangle -> another patch
    p2, _, _ = gpu.template_samples(v, c, f, CAM, poses[:2], 3, z_override=np.array([0.35, 0.35], np.float32))
    assert not np.array_equal(p2, patches[:2])
    # device variant on the net's stream: mesh -> patches -> descriptors
    net = gpu.Net(named="manifold_rgbdnor_16")
    net.fill(4, 1)
    d_v, d_c, d_f = gpu.DeviceArray.from_host(v), gpu.DeviceArray.from_host(c), gpu.DeviceArray.from_host(f)
    d_p = gpu.DeviceArray((n, 64, 64, 7))
    d_desc = gpu.DeviceArray((n, 16))
    lab = np.empty((n, 8), np.float32)
    pp = np.ascontiguousarray(poses.reshape(-1, 12), np.float32)
    gpu.check(gpu.lib().pe_template_samples_dev(net.h, 0, d_v.ptr, d_c.ptr, None, len(v), d_f.ptr, len(f), None, 0, 0, CAM.ctypes.data, pp.ctypes.data, n, 3,
                                                None, 480, 640, 0.001, 4.0, 0.2, d_p.ptr, lab.ctypes.data, None))
    net.compute_descriptors_dev(d_p, n, d_desc)
    net.sync()
    plain, plain_lab, _ = gpu.template_samples(v, c, f, CAM, poses, 3)
    assert np.array_equal(d_p.to_host(), plain) and np.array_equal(lab, plain_lab)
    assert np.array_equal(d_desc.to_host(), net.compute_descriptors(plain))
