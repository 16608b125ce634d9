"""Patch extraction: the CPU oracle (oracle/oracle_patch.c) against the committed cv2 / numpy pins
(tests/golden/golden_patch.npz, made by tests/golden/make_golden_patch.py).  CPU only."""
import os

import numpy as np
import pytest

from tests import scenes
from tests.golden import make_golden_patch as mgp

GOLD = os.path.join(os.path.dirname(__file__), "golden", "golden_patch.npz")


@pytest.fixture(scope="module")
def gp():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def frame():
    color, depth, _ = scenes.make_frame(mgp.SEED, mgp.ROWS, mgp.COLS)
    cam = scenes.cam_for(mgp.ROWS, mgp.COLS)
    cam["fy"] = cam["fx"]
    return color, depth, cam


def test_depth2normals_matches_second_transcription(po, gp, frame):
    color, depth, cam = frame
    nor = po.depth2normals(depth, cam["fx"], cam["fy"], cam["ox"], cam["oy"])
    want = gp["normals_even"]
    got = nor[::2, ::2]
    assert np.array_equal(got == 0, want == 0)                       # same undefined pixels (border, holes, degenerate)
    np.testing.assert_allclose(got, want, rtol=0, atol=2e-6)
    lens = np.linalg.norm(nor, axis=-1)
    assert np.all((np.abs(lens - 1) < 1e-6) | (lens == 0))
    assert np.all(nor[:3] == 0) and np.all(nor[-4:] == 0) and np.all(nor[:, :3] == 0) and np.all(nor[:, -4:] == 0)


def test_sample_patch_matches_cv2(po, gp, frame):
    color, depth, cam = frame
    nor = po.depth2normals(depth, cam["fx"], cam["fy"])
    for i, (cx, cy, z) in enumerate(mgp.cases(cam)):
        ok, rect = po.patch_rect(mgp.ROWS, mgp.COLS, cx, cy, z, cam["fx"], cam["fy"])
        assert ok and rect == tuple(gp["rects"][i])
        got = po.sample_patch(color, depth, nor, cx, cy, z, cam["fx"], cam["fy"])
        want = gp[f"patch_{i}"]
        assert np.array_equal(got[..., 3], want[..., 3])             # nearest-neighbour depth: same pixels, same arithmetic
        np.testing.assert_allclose(got, want, rtol=0, atol=3e-7)     # bilinear channels: cv2 contracts to FMA, 1 ulp
    assert tuple(gp["rects"][-1][2:]) == (128, 128)                  # the INTER_AREA reroute is among the cases


def test_sample_patch_rejects_what_the_reference_asserts(po, frame):
    color, depth, cam = frame
    nor = np.zeros(depth.shape + (3,), np.float32)
    assert po.sample_patch(color, depth, nor, 100, 100, 0.1, cam["fx"], cam["fy"]) is None      # cut larger than the frame
    ok, _ = po.patch_rect(mgp.ROWS, mgp.COLS, 100, 100, 1e9, cam["fx"], cam["fy"])
    assert not ok                                                                                # zero-sized cut
