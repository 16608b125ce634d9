#!/usr/bin/env python
"""Throughput of the template pass on the device (SURVEY 8f rank 1 + 2 + the forward): the view sphere of one object
(depth 2, 7 in-plane rotations, 3 distances = 1869 views) rendered at 640x480 from a 20,480-triangle mesh, one patch per
view, one descriptor per patch.  (CPU timings of the oracle renderer: tests/ only - the oracle is test infrastructure.)"""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
capi = importlib.import_module("3dposeestimation_b200.capi")
from tests import scenes

CAM = np.array([572.4114, 573.57043, 325.2611, 242.04899], np.float32)
LEVEL = int(os.environ.get("LEVEL", 5)); B = int(os.environ.get("BATCH", 128))
v, f = scenes.icosphere(LEVEL, 0.09)
v = (v * (1 + 0.1 * np.sin(9 * v[:, :1]))).astype(np.float32)
c = (v / 0.2 + 0.5).clip(0, 1).astype(np.float32)
poses = capi.sphere_views(2, (0.5, 0.2, 0.9), (-45, 15, 45))
n = len(poses)
net = capi.Net(named="manifold_rgbdnor_16"); net.fill(1, 1)
d_v, d_c, d_f = capi.DeviceArray.from_host(v), capi.DeviceArray.from_host(c), capi.DeviceArray.from_host(f)
d_pose = capi.DeviceArray.from_host(poses)
d_col = capi.DeviceArray((B, 480, 640, 3), np.uint8); d_dep = capi.DeviceArray((B, 480, 640))
ctr = np.tile(np.array([[int(CAM[2]), int(CAM[3])]], np.int32), (n, 1)); z = poses[:, 2, 3].copy()
d_ctr, d_z = capi.DeviceArray.from_host(ctr), capi.DeviceArray.from_host(z)
d_p = capi.DeviceArray((B, 64, 64, 7)); d_desc = capi.DeviceArray((n, 16))

def run(render_only):
    for first in range(0, n, B):
        m = min(B, n - first)
        capi.check(capi.lib().pe_render_views_dev(net.h, 0, d_v.ptr, d_c.ptr, len(v), d_f.ptr, len(f), CAM.ctypes.data, d_pose.ptr + first * 48, m,
                                                  480, 640, 0.001, 4.0, d_col.ptr, d_dep.ptr))
        if render_only:
            continue
        capi.check(capi.lib().pe_sample_patches_dev(net.h, 0, d_col.ptr, d_dep.ptr, None, m, 480, 640, None, d_ctr.ptr + first * 8, d_z.ptr + first * 4,
                                                    m, float(CAM[0]), float(CAM[1]), 0.2, d_p.ptr, None))
        net.compute_descriptors_dev(d_p, m, d_desc, out_offset=first)

def timed(fn, reps=3):
    fn(); net.sync(); t = capi.Timer(0); t.start(net)
    for _ in range(reps): fn()
    t.stop(net); return t.elapsed_ms() / reps

ms_r, ms_all = timed(lambda: run(True)), timed(lambda: run(False))
print(json.dumps({"views": n, "triangles": int(len(f)), "vertices": int(len(v)), "batch": B, "ms_render_all_views": ms_r, "views_per_s_render": n / ms_r * 1e3,
                  "Mtri_per_s": n * len(f) / ms_r / 1e3, "ms_render_patch_descriptor": ms_all, "templates_per_s": n / ms_all * 1e3}))
