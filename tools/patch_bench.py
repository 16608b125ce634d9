#!/usr/bin/env python
"""Throughput of the patch extraction on device-resident 640x480 frames: one 64x64x7 patch per frame
(the reference's use), normals on the fly vs. full-frame depth2normals first."""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
capi = importlib.import_module("3dposeestimation_b200.capi")
from tests import scenes

F = int(os.environ.get("FRAMES", 1024))
base = [scenes.make_frame(s) for s in range(8)]
color = np.stack([base[i % 8][0] for i in range(F)])
depth = np.stack([base[i % 8][1] for i in range(F)])
ctr = np.array([base[i % 8][2] for i in range(F)], np.int32)
z = np.array([depth[i, ctr[i, 1], ctr[i, 0]] for i in range(F)], np.float32)
cam = scenes.CAM
net = capi.Net(named="manifold_rgbdnor_16"); net.fill(1, 1)
d_col, d_dep = capi.DeviceArray.from_host(color), capi.DeviceArray.from_host(depth)
d_ctr, d_z = capi.DeviceArray.from_host(ctr), capi.DeviceArray.from_host(z)
d_nor = capi.DeviceArray((F, 480, 640, 3))
d_p = capi.DeviceArray((F, 64, 64, 7)); d_desc = capi.DeviceArray((F, 16))

def timed(fn, reps=10):
    for _ in range(2): fn()
    net.sync(); t = capi.Timer(0); t.start(net)
    for _ in range(reps): fn()
    t.stop(net); return t.elapsed_ms() / reps

fly = lambda: capi.sample_patches_dev(net, d_col, d_dep, None, F, 480, 640, None, d_ctr, d_z, F, cam["fx"], cam["fy"], 0.2, d_p)
full = lambda: (capi.check(capi.lib().pe_depth2normals_dev(net.h, 0, d_dep.ptr, F, 480, 640, cam["fx"], cam["fy"], 0.0, 0.0, d_nor.ptr)),
                capi.sample_patches_dev(net, d_col, d_dep, d_nor, F, 480, 640, None, d_ctr, d_z, F, cam["fx"], cam["fy"], 0.2, d_p))
chain = lambda: (fly(), net.compute_descriptors_dev(d_p, F, d_desc))
ms_fly, ms_full, ms_chain = timed(fly), timed(full), timed(chain)
# the cut rectangle of samplePatchWithScale (reference src/datasetmanager.cpp:303-309): 0.2 m scaled by f / z
w, h = int(cam["fx"] * 0.2 / float(z[0])), int(cam["fy"] * 0.2 / float(z[0]))
bytes_patch = 64 * 64 * 7 * 4 + (w + 8) * (h + 8) * 7          # output + the cut (colour 3 B, depth 4 B) with the stencil halo
print(json.dumps({"frames": F, "cut": [w, h], "ms_sample_normals_on_the_fly": ms_fly, "patches_per_s": F / ms_fly * 1e3,
                  "algorithmic_GBps": F * bytes_patch / ms_fly / 1e6, "ms_depth2normals_full_plus_sample": ms_full,
                  "full_frame_normals_GBps": F * 480 * 640 * 16 / (ms_full - ms_fly + 1e-9) / 1e6,
                  "ms_frames_to_descriptors": ms_chain}))
