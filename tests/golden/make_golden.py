#!/usr/bin/env python
"""Generates tests/golden/golden.npz - the pins of the CPU oracle.

The reference has no tests, fixtures or golden vectors and its arithmetic lives in
Caffe / OpenCV-C++ / Eigen, none of which can be built in this container.  This
script pins the oracle against INDEPENDENT implementations that are importable here:

  forward    torch (CPU, fp32 and fp64): conv2d, max_pool2d(ceil_mode=True), relu/elu/prelu, linear
  L2 kNN     cv2.BFMatcher(cv2.NORM_L2).knnMatch          (the Python face of the reference's matcher)
  Hamming    cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch
  angles     scipy.spatial.transform.Rotation (float64)
  histogram / confusion / bootstrap: a second, separately written numpy transcription of
             reference src/networkevaluator.cpp:199-355 and src/networksolver.cpp:474-504

Inputs are regenerated from the splitmix64 generators in 3dposeestimation_b200/synth.py, so only
outputs are stored.  Run from the repo root:  python tests/golden/make_golden.py
"""
import importlib
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
synth = importlib.import_module("3dposeestimation_b200.synth")
from oracle import pyoracle as po  # noqa: E402  (prototxt reader only; no oracle arithmetic here)
from tests import netdefs  # noqa: E402

SEED = synth.BASE_SEED


# ------------------------------------------------------------------------------ forward ---
def torch_forward(net, blobs, patches, dtype):
    import torch
    import torch.nn.functional as F
    torch.set_num_threads(1)
    x = torch.from_numpy(patches).to(dtype).permute(0, 3, 1, 2)[:, net.sel]  # HWC -> CHW + select
    stages = []
    for kind, _, a, b, b0, b1 in net.layers:
        if kind == po.CONV:
            x = F.conv2d(x, torch.from_numpy(blobs[b0]).to(dtype), torch.from_numpy(blobs[b1]).to(dtype))
        elif kind == po.POOL:
            x = F.max_pool2d(x, a, b, ceil_mode=True)
        elif kind == po.ACT:
            if a == 1:
                x = F.relu(x)
            elif a == 2:
                x = F.elu(x, alpha=1.0)
            else:
                x = F.prelu(x, torch.from_numpy(blobs[b0]).to(dtype))
        else:
            x = F.linear(x.flatten(1), torch.from_numpy(blobs[b0]).to(dtype), torch.from_numpy(blobs[b1]).to(dtype))
        stages.append(x.detach().numpy().copy())
    return x.detach().numpy(), stages


# ----------------------------------------------------------------- independent glue code ---
def ang_dist64(qa, qb):
    d = abs(float(np.dot(qa.astype(np.float64), qb.astype(np.float64))))
    return 2.0 * math.acos(min(1.0, d))


def py_histogram(knn_idx, ql, tl, nr_tmpl, rot_inv, bins):
    """float32 re-transcription of computeHistogram's loop (src/networkevaluator.cpp:269-355)"""
    f32 = np.float32
    histo = np.zeros(len(bins), np.float32)
    mean = f32(0)
    diffs = []

    def ang32(a, b):
        d = f32(0)
        for i in range(4):
            d = f32(d + f32(a[i] * b[i]))
        d = min(f32(1), abs(d))
        return f32(f32(2) * np.arccos(f32(d)))

    def r22(q):
        tx, ty = f32(2) * q[0], f32(2) * q[1]
        return f32(f32(1) - f32(f32(tx * q[0]) + f32(ty * q[1])))

    def deg(x):
        return f32(np.float64(f32(x * f32(180))) / math.pi)

    for qi in range(knn_idx.shape[0]):
        qq = ql[qi, 1:5]
        best, best_t = f32(np.finfo(np.float32).max), -1
        for nn in range(knn_idx.shape[1]):
            t = int(knn_idx[qi, nn])
            if ql[qi, 0] != tl[t, 0]:
                continue
            if rot_inv[int(ql[qi, 0])] == 0:
                d = deg(ang32(qq, tl[t, 1:5]))
            else:
                with np.errstate(invalid="ignore"):
                    d = deg(abs(f32(np.arccos(r22(qq)) - np.arccos(r22(tl[t, 1:5])))))
                if np.isnan(d):
                    d = f32(np.finfo(np.float32).max)
            if d >= best:
                continue
            best, best_t = d, t
        if best_t < 0:
            histo[0] += 1
            continue
        mean = f32(mean + best)
        diffs.append(best)
        obj = best_t // nr_tmpl
        ref = ang32(qq, tl[best_t, 1:5])
        exact = True
        for p in range(nr_tmpl):
            if ang32(qq, tl[obj * nr_tmpl + p, 1:5]) < ref:
                exact = False
        if exact:
            histo[1] += 1
        for b in range(2, len(bins)):
            if best < bins[b]:
                histo[b] += 1
    total = f32(histo[0] + histo[-1])
    histo = (histo / total).astype(np.float32)
    mean = f32(mean / f32(knn_idx.shape[0]))
    diffs.sort()
    med = diffs[(len(diffs) - 1) // 2] if diffs else np.float32("nan")
    return np.concatenate([histo, [mean, med]]).astype(np.float32)


def py_confusion(knn_idx, ql, tl, c2l, n_local, nr_test):
    cm = np.zeros((n_local, n_local), np.float32)
    for qi in range(knn_idx.shape[0]):
        k = knn_idx.shape[1]
        for nn in range(k):
            t = int(knn_idx[qi, nn])
            if ql[qi, 0] == tl[t, 0]:
                cm[c2l[int(ql[qi, 0])], c2l[int(tl[t, 0])]] += 1
                break
            if nn == k - 1:
                t0 = int(knn_idx[qi, 0])
                cm[c2l[int(ql[qi, 0])], c2l[int(tl[t0, 0])]] += 1
    return cm / np.float32(nr_test)


def py_bootstrap(knn3, nr_q, nr_t):
    out = -np.ones((knn3.shape[0], 4), np.int32)
    ln = np.zeros(knn3.shape[0], np.int32)
    for qi in range(knn3.shape[0]):
        qobj = qi // nr_q
        lst = []
        for nn in range(3):
            o, p = int(knn3[qi, nn]) // nr_t, int(knn3[qi, nn]) % nr_t
            if nn == 0:
                lst += [o, p]
            elif len(lst) < 4 and o != qobj:
                lst += [o, p]
        out[qi, :len(lst)] = lst
        ln[qi] = len(lst)
    return out, ln


def cv2_knn(q, t, k, norm):
    import cv2
    m = cv2.BFMatcher(norm)
    m.add([t])
    res = m.knnMatch(q, k)
    idx = np.array([[d.trainIdx for d in r] for r in res], np.int32)
    dist = np.array([[d.distance for d in r] for r in res], np.float32)
    return idx, dist


def knn_problem(seed, nq, nt, dim):
    """class-free random descriptors for the matcher goldens"""
    t = synth.uniform(seed, 30, (nt, dim), -1.0, 1.0)
    q = synth.uniform(seed, 31, (nq, dim), -1.0, 1.0)
    return q, t


def eval_problem(seed, n_obj, nr_t, nr_q, dim):
    """descriptors with class structure: class centre + pose offset + noise"""
    centre = 0.5 * synth.uniform(seed, 40, (n_obj, 1, dim), -1.0, 1.0)
    t = centre + 0.6 * synth.uniform(seed, 41, (n_obj, nr_t, dim), -1.0, 1.0)
    src = (np.arange(nr_q) * 7) % nr_t
    q = t[:, src, :] + 0.8 * synth.uniform(seed, 42, (n_obj, nr_q, dim), -1.0, 1.0)
    tl = synth.make_labels(seed, n_obj, nr_t)
    ql = synth.perturb_labels(tl.reshape(n_obj, nr_t, 8)[:, src].reshape(-1, 8), seed)
    return (q.reshape(-1, dim).astype(np.float32), t.reshape(-1, dim).astype(np.float32), ql, tl)


def main():
    import cv2
    from scipy.spatial.transform import Rotation as R
    g = {}
    # 1. forward of every variant
    for vi, name in enumerate(netdefs.VARIANTS):
        net = po.parse_prototxt(netdefs.prototxt(name))
        blobs = synth.make_blobs(net.blob_names, net.blob_shapes, SEED + vi, mode=1)
        patches = synth.make_patches(SEED, 100 * vi, 3, net.channels, 2)
        d32, st32 = torch_forward(net, blobs, patches, __import__("torch").float32)
        d64, _ = torch_forward(net, blobs, patches, __import__("torch").float64)
        g[f"fwd32_{name}"] = d32.astype(np.float32)
        g[f"fwd64_{name}"] = d64.astype(np.float64)
        if name == "manifold_rgbdnor_16":
            g["stage2_pool1act"] = st32[2][0].astype(np.float32)   # 16x29x29 of patch 0
            g["stage5_pool2act"] = st32[5][0].astype(np.float32)   # 7x13x13
        print(name, d32[0, :3], float(np.abs(d32 - d64).max()))
    # 2. L2 matcher (incl. duplicated rows -> ties)
    for dim in (16, 3, 64):
        q, t = knn_problem(SEED + dim, 64, 512, dim)
        t[100] = t[7]
        t[300] = t[7]
        q[5] = t[7]
        idx, dist = cv2_knn(q, t, 5, cv2.NORM_L2)
        g[f"knn_idx_{dim}"], g[f"knn_dist_{dim}"] = idx, dist
    # 3. Hamming matcher
    codes_t = (synth.uniform01(SEED, 50, 512 * 4) * 2 ** 24).astype(np.uint32).reshape(512, 4)
    codes_t = codes_t * np.uint32(257) ^ (codes_t >> np.uint32(3))
    codes_q = codes_t[(np.arange(64) * 5) % 512] ^ (np.uint32(1) << (np.arange(64, dtype=np.uint32) % 32))[:, None]
    hi, hd = cv2_knn(codes_q.view(np.uint8).reshape(64, 16), codes_t.view(np.uint8).reshape(512, 16), 5,
                     cv2.NORM_HAMMING)
    g["ham_idx"], g["ham_dist"] = hi, hd.astype(np.int32)
    # 4. quaternion maths
    qa = synth.random_quaternions(SEED, 60, 256)
    qb = synth.random_quaternions(SEED, 61, 256)
    qb[:8] = qa[:8]            # identical -> 0
    qb[8:16] = -qa[8:16]       # antipodal -> 0
    rel = R.from_quat(qa.astype(np.float64)).inv() * R.from_quat(qb.astype(np.float64))
    g["angles_rad"] = rel.magnitude()
    g["rot22"] = R.from_quat(qa.astype(np.float64)).as_matrix()[:, 2, 2]
    # 5. histogram / confusion / bootstrap on cv2 neighbours
    n_obj, nr_t, nr_q, dim = 6, 40, 15, 16
    q, t, ql, tl = eval_problem(SEED, n_obj, nr_t, nr_q, dim)
    rot_inv = np.array([0, 0, 1, 0, 2, 0], np.int32)
    bins = np.array([-1, 0, 10, 20, 40, 180], np.float32)
    for k in (1, 3, 5):
        idx, _ = cv2_knn(q, t, k, cv2.NORM_L2)
        g[f"eval_idx_k{k}"] = idx
        g[f"eval_hist_k{k}"] = py_histogram(idx, ql, tl, nr_t, rot_inv, bins)
        c2l = np.arange(n_obj, dtype=np.int32)[::-1].copy()
        g[f"eval_conf_k{k}"] = py_confusion(idx, ql, tl, c2l, n_obj, nr_q)
        print("hist k", k, g[f"eval_hist_k{k}"])
    bo, bl = py_bootstrap(g["eval_idx_k3"], nr_q, nr_t)
    g["eval_boot"], g["eval_boot_len"] = bo, bl
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "golden.npz"), **g)
    print("wrote golden.npz with", len(g), "arrays")


if __name__ == "__main__":
    main()
