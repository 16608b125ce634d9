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
