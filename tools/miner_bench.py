#!/usr/bin/env python
"""Throughput of the triplet miner at LineMOD scale: 15 objects x 1,241 templates x 1,200 training
samples of 64x64x7 (both pools resident in HBM, filled by the device generator): computeMaxSimTmpl, computeKNN
(forward of both pools + 3-NN + table) and buildBatch (selection + Caffe-layout gather [+ colour fill]) for the
reference's batch of 300 and for a 30,000-sample batch."""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
capi = importlib.import_module("3dposeestimation_b200.capi")
from tests import miner_cases as mc

N_OBJ, N_T, N_Q, CH = 15, int(os.environ.get("TEMPLATES", 1241)), int(os.environ.get("TRAINING", 1200)), 7
tl, ql = mc.make_labels(1, N_OBJ, N_T), mc.make_labels(2, N_OBJ, N_Q)
rot_inv = np.array([0, 0, 1, 0, 0, 0, 2, 0, 0, 2, 2, 0, 0, 0, 0], np.int32)       # config.ini:10
net = capi.Net(named="manifold_rgbdnor_16"); net.fill(1, 1)
m = capi.Miner(None, tl, None, ql, N_OBJ, rot_inv, inplane=True, channels=CH)
m.share_stream(net)
d_t, d_q = m.pools()
for first in range(0, N_OBJ * N_T, 4096):
    n = min(4096, N_OBJ * N_T - first)
    net.synth_patches_dev(7, first, n, N_T, d_t + first * 4096 * CH * 4)
for first in range(0, N_OBJ * N_Q, 4096):
    n = min(4096, N_OBJ * N_Q - first)
    net.synth_patches_dev(8, first, n, N_Q, d_q + first * 4096 * CH * 4)
net.sync()

def timed(fn, reps=10, warm=2):
    for _ in range(warm): fn()
    net.sync(); t = capi.Timer(0); t.start(net)
    for _ in range(reps): fn()
    t.stop(net); return t.elapsed_ms() / reps

t0 = time.perf_counter(); ms_tab = m.max_sim(); ms_maxsim_first = (time.perf_counter() - t0) * 1e3
t0 = time.perf_counter(); m.compute_knn(net); ms_knn_first = (time.perf_counter() - t0) * 1e3
t0 = time.perf_counter(); table, length = m.compute_knn(net); ms_knn = (time.perf_counter() - t0) * 1e3
out = {"objects": N_OBJ, "templates_per_object": N_T, "training_per_object": N_Q, "channels": CH,
       "ms_max_sim_first_call": ms_maxsim_first, "ms_compute_knn_first_call": ms_knn_first, "ms_compute_knn": ms_knn,
       "compute_knn_patches_per_s": N_OBJ * (N_T + N_Q) / ms_knn * 1e3,
       "knn_hits_pct": float((table[:, 1] == ms_tab[:, 0]).mean() * 100)}
img = 64 * 64 * CH * 4
for batch in (300, 30000):
    d_b = capi.DeviceArray((batch, CH, 64, 64))
    state = {"it": 0}
    def step(fill=0):
        m.build_batch_dev(batch, state["it"] % (N_OBJ * N_Q * 5 // batch), 11 + state["it"], d_b, random_background=fill)
        state["it"] += 1
    ms0 = timed(step, reps=50 if batch == 300 else 10)
    assert not m.flags(batch // 5).any()
    ms1 = timed(lambda: step(1), reps=50 if batch == 300 else 10)
    out[f"batch_{batch}"] = {"ms_build_batch": ms0, "samples_per_s": batch / ms0 * 1e3, "algorithmic_GBps": 2 * batch * img / ms0 / 1e6,
                             "ms_with_colour_fill": ms1}
    d_b.free()
t = capi.Timer(0); m2 = capi.Miner(None, tl, None, ql, N_OBJ, rot_inv, inplane=True, channels=CH); m2.share_stream(net)
net.sync(); t.start(net); m2.max_sim(); t.stop(net); out["ms_max_sim_kernel"] = t.elapsed_ms()
print(json.dumps(out))
