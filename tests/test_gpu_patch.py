"""GPU parity of the patch extraction through the C ABI: pe_depth2normals / pe_sample_patches against
the CPU oracle and the committed cv2 pins, then frames -> patches -> descriptors on the device."""
import os

import numpy as np
import pytest

from tests import netdefs, scenes
from tests.golden import make_golden_patch as mgp

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "golden_patch.npz")


def test_depth2normals_bit_exact_vs_oracle(gpu, po):
    frames = [scenes.make_frame(s, 480, 640) for s in (1, 2)]
    depth = np.stack([f[1] for f in frames])
    cam = scenes.CAM
    got = gpu.depth2normals(depth, cam["fx"], cam["fy"], cam["ox"], cam["oy"])
    for i in range(2):
        want = po.depth2normals(depth[i], cam["fx"], cam["fy"], cam["ox"], cam["oy"])
        assert np.array_equal(got[i], want)        # same operations in the same order, no FMA on either side


@pytest.mark.parametrize("precomputed", [False, True], ids=["normals-on-the-fly", "normals-given"])
def test_sample_patches_vs_oracle_and_cv2(gpu, po, precomputed):
    gp = np.load(GOLD)
    color, depth, _ = scenes.make_frame(mgp.SEED, mgp.ROWS, mgp.COLS)
    cam = scenes.cam_for(mgp.ROWS, mgp.COLS)
    cam["fy"] = cam["fx"]
    cs = mgp.cases(cam) + [(100, 100, 0.1), (60, 200, 1.37), (300, 20, 0.71)]       # one the reference would assert on
    ctr = np.array([[c[0], c[1]] for c in cs], np.int32)
    z = np.array([c[2] for c in cs], np.float32)
    nor = po.depth2normals(depth, cam["fx"], cam["fy"])
    patches, status = gpu.sample_patches(color[None], depth[None], ctr, z, cam["fx"], cam["fy"], 0.2,
                                         normals=nor[None] if precomputed else None, frame_of=np.zeros(len(cs), np.int32))
    for i, (cx, cy, zz) in enumerate(cs):
        want = po.sample_patch(color, depth, nor, cx, cy, zz, cam["fx"], cam["fy"])
        if want is None:
            assert status[i] == 1 and not patches[i].any()
            continue
        assert status[i] == 0
        assert np.array_equal(patches[i], want), i                    # bit-exact vs the oracle
        if i < 4:
            np.testing.assert_allclose(patches[i], gp[f"patch_{i}"], rtol=0, atol=3e-7)   # and within 1 ulp of cv2


def test_frames_to_descriptors_on_device(gpu, po, synth):
    """camera buffers -> patches -> descriptors without a host round trip == the same chain through host buffers"""
    n = 6
    frames = [scenes.make_frame(10 + s, 480, 640) for s in range(n)]
    color = np.stack([f[0] for f in frames])
    depth = np.stack([f[1] for f in frames])
    ctr = np.array([f[2] for f in frames], np.int32)
    z = np.array([depth[i, ctr[i, 1], ctr[i, 0]] for i in range(n)], np.float32)
    cam = scenes.CAM
    ref = po.parse_prototxt(netdefs.prototxt("manifold_rgbdnor_16"))
    blobs = synth.make_blobs(ref.blob_names, ref.blob_shapes, 3, mode=1)
    net = gpu.Net(text=netdefs.prototxt("manifold_rgbdnor_16"))
    net.set_blobs(blobs)
    d_col = gpu.DeviceArray.from_host(color)
    d_dep = gpu.DeviceArray.from_host(depth)
    d_ctr = gpu.DeviceArray.from_host(ctr)
    d_z = gpu.DeviceArray.from_host(z)
    d_p = gpu.DeviceArray((n, 64, 64, 7))
    d_st = gpu.DeviceArray((n,), np.int32)
    d_desc = gpu.DeviceArray((n, 16))
    gpu.sample_patches_dev(net, d_col, d_dep, None, n, 480, 640, None, d_ctr, d_z, n, cam["fx"], cam["fy"], 0.2, d_p, d_st)
    net.compute_descriptors_dev(d_p, n, d_desc)
    net.sync()
    assert not d_st.to_host().any()
    patches, status = gpu.sample_patches(color, depth, ctr, z, cam["fx"], cam["fy"])
    assert np.array_equal(d_p.to_host(), patches)
    assert np.array_equal(d_desc.to_host(), net.compute_descriptors(patches))
    want = po.OracleNet(ref, blobs).forward(patches)
    scale = np.abs(want).max(axis=-1, keepdims=True)
    assert float((np.abs(d_desc.to_host() - want) / scale).max()) < 1e-4


def test_scene_samples_equal_the_oracle_chain(gpu, po):
    """pe_scene_samples = datasetManager::extractRealSamplesPaul / Wadim (src/datasetmanager.cpp:348-411): projected
    centroid (truncated to int) -> normals -> patch -> label, against the oracle's stages chained by hand"""
    from scipy.spatial.transform import Rotation
    cam = np.array([scenes.CAM["fx"], scenes.CAM["fy"], scenes.CAM["ox"], scenes.CAM["oy"]], np.float32)
    frames = [scenes.make_frame(40 + i) for i in range(4)]
    color = np.stack([f[0] for f in frames])
    depth = np.stack([f[1] for f in frames])
    poses = np.zeros((4, 3, 4), np.float32)
    poses[:, :, :3] = Rotation.random(4, random_state=3).as_matrix().astype(np.float32)
    for i, (_, dep, (cx, cy)) in enumerate(frames):       # ground-truth translation = the sphere's surface point back-projected
        z = float(dep[cy, cx])
        poses[i, :, 3] = [(cx + 0.3 - cam[2]) * z / cam[0], (cy + 0.6 - cam[3]) * z / cam[1], z]
    centroid = np.array([0.01, -0.02, 0.015], np.float32)
    for kw, ora in (({}, dict(centroid=None, from_depth=False)), ({"z_from_depth": True}, dict(centroid=None, from_depth=True)),
                    ({"centroid": centroid}, dict(centroid=centroid, from_depth=False))):
        patches, labels, status = gpu.scene_samples(color, depth, cam, poses, 5, **kw)
        assert not status.any()
        for i in range(4):
            x, y, z = po.project_centroid(cam, poses[i], ora["centroid"])
            if ora["from_depth"]:
                z = float(depth[i, y, x])
            nor = po.depth2normals(depth[i], float(cam[0]), float(cam[1]))
            want = po.sample_patch(color[i], depth[i], nor, x, y, z, float(cam[0]), float(cam[1]))
            assert np.array_equal(patches[i], want), (kw, i)
            assert np.array_equal(labels[i], po.pose_label(poses[i], 5))
    # the projected centroid of frame 0 is where the scene put it (truncation, not rounding)
    x, y, _ = po.project_centroid(cam, poses[0])
    assert (x, y) == (frames[0][2][0], frames[0][2][1])
    # the Wadim variant refuses a centroid over missing depth / outside the frame (the reference asserts)
    bad = poses.copy()
    bad[0, :, 3] = [(80 - cam[2]) * 1.0 / cam[0], (70 - cam[3]) * 1.0 / cam[1], 1.0]      # inside the sensor-hole rectangle of make_frame
    with pytest.raises(gpu.PoseError) as e:
        gpu.scene_samples(color, depth, cam, bad, 5, z_from_depth=True)
    assert e.value.code == gpu.PE_ERR_INVALID and "depth" in str(e.value)
    bad[0, :, 3] = [5.0, 0.0, 1.0]
    with pytest.raises(gpu.PoseError):
        gpu.scene_samples(color, depth, cam, bad, 5, z_from_depth=True)
