#!/usr/bin/env python
"""Generates tests/golden/golden_patch.npz - pins of the patch-extraction oracle (oracle/oracle_patch.c).

The reference's samplePatchWithScale (src/datasetmanager.cpp:299-346) is OpenCV calls around a few scalar lines; the pins
come from the same calls through cv2 (4.13 here; the reference names 3.1 - the INTER_LINEAR / INTER_NEAREST / 2x -> INTER_AREA
rules of resize() are the same): crop, convertTo(1/255), the depth de-mean / clamp / rescale written with numpy float32
ops, cv2.resize to 64 x 64, channel merge.  depth2normals (src/helper.cpp:53-103) has no library twin: it is pinned by a
second, vectorised numpy float32 transcription.  Inputs are regenerated from tests/scenes.py; only outputs are stored.
Run from the repo root:  python tests/golden/make_golden_patch.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import scenes  # noqa: E402

ROWS, COLS, SEED = 240, 320, 77
# (center_x, center_y, z): plain, near + clamped at the top-left corner, far (up-scaling) + clamped at the bottom-right
# corner, and a cut of exactly 128 px
CASES = [(150, 118, 0.9), (3, 5, 0.45), (318, 237, 2.6), (140, 120, None)]


def np_depth2normals(dep, fx, fy):
    """vectorised float32 transcription of src/helper.cpp:53-103 (accumulation order i outer, j inner)"""
    f32 = np.float32
    rows, cols = dep.shape
    N = 3
    nor = np.zeros((rows, cols, 3), f32)
    rs, cs = slice(N, rows - N - 1), slice(N, cols - N - 1)
    d = dep[rs, cs]
    A0 = np.zeros_like(d); A1 = np.zeros_like(d); A3 = np.zeros_like(d); b0 = np.zeros_like(d); b1 = np.zeros_like(d)
    for i in (-N, 0, N):
        for j in (-N, 0, N):
            nb = dep[N + j:rows - N - 1 + j, N + i:cols - N - 1 + i]
            delta = (nb - d).astype(f32)
            f = (np.abs(delta) < f32(0.05)).astype(f32)
            fi, fj = f * f32(i), f * f32(j)
            A0 = A0 + fi * f32(i); A1 = A1 + fi * f32(j); A3 = A3 + fj * f32(j)
            b0 = b0 + fi * delta; b1 = b1 + fj * delta
    n0 = (A3 * b0 - A1 * b1) * f32(fx)
    n1 = (-A1 * b0 + A0 * b1) * f32(fy)
    n2 = (A0 * A3 - A1 * A1) * d
    sq = n0 * n0 + n1 * n1 + n2 * n2
    ok = (d > 0) & (sq > 0)
    s = np.sqrt(np.where(ok, sq, 1)).astype(f32)
    out = np.stack([n0 / s, n1 / s, n2 / s], -1).astype(f32)
    nor[rs, cs] = np.where(ok[..., None], out, 0)
    return nor


def cut_rect(rows, cols, cx, cy, z, fx, fy, m):
    f32 = np.float32
    w, h = int(f32(fx) * f32(m) / f32(z)), int(f32(fy) * f32(m) / f32(z))
    x, y = cx - w // 2, cy - h // 2          # C++ int division of non-negative w, h
    x, y = max(x, 0), max(y, 0)
    x, y = min(x, cols - w - 1), min(y, rows - h - 1)
    return x, y, w, h


def cv2_patch(color, depth, normals, cx, cy, z, fx, fy, m=0.2):
    import cv2
    f32 = np.float32
    x, y, w, h = cut_rect(depth.shape[0], depth.shape[1], cx, cy, z, fx, fy, m)
    col = color[y:y + h, x:x + w].astype(f32) * (f32(1) / f32(255))
    dep = depth[y:y + h, x:x + w].copy()
    nor = normals[y:y + h, x:x + w].copy()
    dep -= f32(z)
    dep[dep < -f32(m)] = -f32(m)
    dep[dep > f32(m)] = f32(m)
    dep *= f32(1.0) / f32(m)
    dep = (dep + f32(1)) * f32(0.5)
    col = cv2.resize(col, (64, 64))
    nor = cv2.resize(nor, (64, 64))
    dep = cv2.resize(dep, (64, 64), interpolation=cv2.INTER_NEAREST)
    return np.concatenate([col, dep[..., None], nor], -1).astype(f32), (x, y, w, h)


def cases(cam):
    out = []
    for cx, cy, z in CASES:
        if z is None:                 # the z that makes the cut exactly 128 x 128 -> cv::resize takes its INTER_AREA route
            z = float(np.float32(cam["fx"]) * np.float32(0.2) / np.float32(128.4))
        out.append((cx, cy, float(np.float32(z))))
    return out


def main():
    color, depth, _ = scenes.make_frame(SEED, ROWS, COLS)
    cam = scenes.cam_for(ROWS, COLS)
    cam["fy"] = cam["fx"]             # square cuts so that the 128 x 128 case exists
    normals = np_depth2normals(depth, cam["fx"], cam["fy"])
    g = {"normals_even": normals[::2, ::2].copy()}          # every other pixel keeps the fixture small
    rects = []
    for i, (cx, cy, z) in enumerate(cases(cam)):
        p, r = cv2_patch(color, depth, normals, cx, cy, z, cam["fx"], cam["fy"])
        g[f"patch_{i}"] = p
        rects.append(r)
    g["rects"] = np.array(rects, np.int32)
    assert tuple(g["rects"][-1][2:]) == (128, 128), g["rects"][-1]
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "golden_patch.npz"), **g)
    print("wrote golden_patch.npz:", {k: v.shape for k, v in g.items()})


if __name__ == "__main__":
    main()
