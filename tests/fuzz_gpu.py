"""Randomised shape sweep of the C-ABI matcher / forward against the oracle (not collected by pytest: run it on a GPU box,
`python -m tests.fuzz_gpu [n_cases] [seed]`).  The fixed shapes of tests/test_gpu_parity.py cover the cases the reference
meets; this walks the space between them: dims 1..128, k 1..16, template counts around the 32-row block / 128-row tile /
chunk boundaries, a single query, duplicated rows, scaled and shifted descriptors, odd batch sizes of every net.
Prints one line per failing case and exits non-zero if any failed."""
import sys

import numpy as np

from tests import netdefs
from tests.test_gpu_parity import check_knn


def knn_cases(rng, n):
    edge_t = [1, 2, 31, 32, 33, 127, 128, 129, 255, 257, 4095, 4097, 8191, 12289]
    for i in range(n):
        dim = int(rng.choice([1, 2, 3, 4, 5, 7, 8, 15, 16, 17, 24, 31, 32, 33, 48, 63, 64, 65, 96, 127, 128]))
        k = int(rng.integers(1, 17))
        nt = int(rng.choice(edge_t)) if rng.random() < 0.5 else int(rng.integers(1, 60000))
        nq = int(rng.choice([1, 2, 3, 127, 128, 129, 255, 256, 257, 600]))
        scale = float(rng.choice([1e-3, 1.0, 1.0, 1.0, 50.0]))
        shift = float(rng.choice([0.0, 0.0, 0.0, 3.0]))
        yield dict(dim=dim, k=k, nt=nt, nq=nq, scale=scale, shift=shift, seed=int(rng.integers(1 << 30)), dup=rng.random() < 0.3)


def run_knn(gpu, po, c):
    rng = np.random.default_rng(c["seed"])
    t = (c["shift"] + c["scale"] * rng.standard_normal((c["nt"], c["dim"]))).astype(np.float32)
    pick = rng.integers(0, c["nt"], c["nq"])
    q = (t[pick] + 0.3 * c["scale"] * rng.standard_normal((c["nq"], c["dim"]))).astype(np.float32)
    if c["dup"] and c["nt"] > 4:
        t[c["nt"] // 2] = t[1]
        t[c["nt"] - 1] = t[1]
        q[0] = t[1]
    db = gpu.TemplateDB(desc=t)
    idx, dist = db.knn(q, c["k"])
    oi, od, gap = po.knn_l2(q, t, c["k"], want_gap=True)
    kk = min(c["k"], c["nt"])
    if c["k"] > c["nt"]:
        assert np.all(idx[:, c["nt"]:] == -1) and np.all(np.isinf(dist[:, c["nt"]:])), "padding"
        assert np.array_equal(oi[:, c["nt"]:], idx[:, c["nt"]:])
    assert np.all(np.diff(dist[:, :kk], axis=1) >= 0), "not ascending"
    assert np.all((idx[:, :kk] >= 0) & (idx[:, :kk] < c["nt"])), "index range"
    for r in range(c["nq"]):
        assert len(set(idx[r, :kk].tolist())) == kk, "duplicate index in a row"
    frac = check_knn(idx[:, :kk], dist[:, :kk], oi[:, :kk], od[:, :kk], gap)
    assert frac <= max(0.02, 1.5 / c["nq"]), f"{frac:.3f} of the rows differ beyond near-ties"
    if c["dup"] and c["nt"] > 4 and kk >= 3:
        assert list(idx[0, :3]) == [1, c["nt"] // 2, c["nt"] - 1] and not dist[0, :3].any(), "tie order"


def run_hamming(gpu, po, rng):
    words = int(rng.choice([1, 2, 4]))
    nt = int(rng.choice([1, 5, 31, 32, 33, 128, 129, 1000, 4097, 20011]))
    nq = int(rng.choice([1, 7, 128, 129, 300]))
    k = int(rng.integers(1, 17))
    t = rng.integers(0, 1 << 32, (nt, words), dtype=np.uint64).astype(np.uint32)
    q = t[rng.integers(0, nt, nq)] ^ (np.uint32(1) << rng.integers(0, 32, (nq, words)).astype(np.uint32))
    if nt > 3:
        t[nt - 1] = t[0]
    idx, dist = gpu.TemplateDB(codes=t).knn_hamming(q, k)
    oi, od = po.knn_hamming(q, t, k)
    assert np.array_equal(idx, oi) and np.array_equal(dist, od), f"hamming words={words} nt={nt} nq={nq} k={k}"
    return dict(words=words, nt=nt, nq=nq, k=k)


def run_forward(gpu, po, synth, rng, name):
    from tests.test_gpu_parity import make_net
    net, oracle, ref = make_net(gpu, po, synth, name, int(rng.integers(1 << 20)))
    n = int(rng.choice([1, 2, 63, 64, 65, 127, 129, 299, 300, 301, 333]))
    x = synth.make_patches(int(rng.integers(1 << 20)), 0, n, ref.channels, 10)
    got = net.compute_descriptors(x)
    want = oracle.forward(x)
    err = np.abs(got - want).max() / (np.abs(want).max() + 1e-30)
    assert got.shape == want.shape and err < 1e-4, f"{name} n={n} err {err:.2e}"
    # a batch gives the same rows as its parts
    if n > 2:
        assert np.array_equal(net.compute_descriptors(x[: n // 2]), got[: n // 2]), f"{name} n={n}: batch-size dependence"
    return dict(net=name, n=n, err=float(err))


def run_eval(gpu, po, synth, rng):
    """computeHistogram / computeConfusionMatrix / computeKNN / computeKNNAccuracy bodies on random neighbour tables"""
    bins = np.array([-1, 0, 10, 20, 40, 180], np.float32)
    n_obj = int(rng.choice([1, 2, 3, 7, 33, 150]))
    nr_t = int(rng.choice([1, 2, 5, 64, 333, 1025]))
    nr_q = int(rng.choice([1, 2, 9, 100, 257]))
    k = int(rng.integers(1, 9))
    seed = int(rng.integers(1 << 20))
    tl = synth.make_labels(seed, n_obj, nr_t)
    src = (np.arange(nr_q) * 13) % nr_t
    ql = synth.perturb_labels(tl.reshape(n_obj, nr_t, 8)[:, src].reshape(-1, 8), seed + 1, max_deg=float(rng.choice([0.0, 5.0, 60.0])))
    rot_inv = rng.integers(0, 3, n_obj).astype(np.int32)
    nq = n_obj * nr_q
    own = (np.repeat(np.arange(n_obj), nr_q) * nr_t)[:, None] + rng.integers(0, nr_t, (nq, k))
    anyc = rng.integers(0, n_obj * nr_t, (nq, k))
    idx = np.where(rng.random((nq, k)) < rng.choice([0.0, 0.5, 1.0]), own, anyc).astype(np.int32)
    if rng.random() < 0.5:
        idx[::3, 0] = (np.repeat(np.arange(n_obj), nr_q) * nr_t + np.tile(src, n_obj))[::3]
    tag = f"eval n_obj={n_obj} nr_t={nr_t} nr_q={nr_q} k={k} seed={seed}"
    db = gpu.TemplateDB(desc=np.zeros((min(4, n_obj * nr_t), 16), np.float32), labels_all=tl, poses_per_object=nr_t, t_total=n_obj * nr_t)
    h = db.histogram(idx, ql, rot_inv, bins)
    ho = po.histogram(idx, ql, tl, nr_t, rot_inv, bins)
    assert np.array_equal(h[:6], ho[:6]), tag + f" bins {h[:6]} vs {ho[:6]}"
    assert np.allclose(h[6:], ho[6:], rtol=0, atol=1e-3, equal_nan=True), tag + f" mean/median {h[6:]} vs {ho[6:]}"
    c2l = rng.permutation(n_obj).astype(np.int32)
    assert np.array_equal(db.confusion(idx, ql, c2l, n_obj, nr_q), po.confusion(idx, ql, tl, c2l, n_obj, nr_q)), tag + " confusion"
    if k >= 3:
        out, ln = db.knn_bootstrap(np.ascontiguousarray(idx[:, :3]), nr_q)
        oo, ol = po.knn_bootstrap(np.ascontiguousarray(idx[:, :3]), nr_q, nr_t)
        assert np.array_equal(out, oo) and np.array_equal(ln, ol), tag + " bootstrap"
        ms = rng.integers(0, nr_t, nq).astype(np.int32)
        ms[::2] = out[::2, 1]
        assert db.knn_accuracy(ms, out, nr_q) == po.knn_accuracy(ms, out, nr_q), tag + " accuracy"


def _rand_pose(rng, dist):
    a = rng.standard_normal(4)
    w, x, y, z = a / np.linalg.norm(a)
    r = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    t = np.array([rng.normal(0, 0.05), rng.normal(0, 0.05), dist])
    return np.concatenate([r, t[:, None]], 1).astype(np.float32)


def run_render(gpu, po, rng):
    """triangle soups the fixed scenes do not hold: slivers, zero-area and repeated-vertex faces, huge triangles, faces behind
    the camera / through the near plane / off screen, odd frame sizes, textured or not - CUDA == oracle bit for bit"""
    from tests import scenes
    kind = int(rng.integers(0, 4))
    if kind == 0:                                   # random soup around the origin
        nv, nf = int(rng.integers(3, 200)), int(rng.integers(1, 300))
        v = (rng.standard_normal((nv, 3)) * rng.choice([0.02, 0.1, 0.6])).astype(np.float32)
        f = rng.integers(0, nv, (nf, 3)).astype(np.int32)            # repeated vertices -> degenerate faces among them
    elif kind == 1:                                 # slivers and coplanar overlaps (z fights resolved by face id)
        nv = 60
        v = np.zeros((nv, 3), np.float32)
        v[:, 0] = rng.uniform(-0.1, 0.1, nv)
        v[:, 1] = rng.uniform(-0.1, 0.1, nv) * rng.choice([1.0, 1e-3])
        f = rng.integers(0, nv, (80, 3)).astype(np.int32)
    elif kind == 2:                                 # a mesh the camera may sit inside
        v, f = scenes.icosphere(2, float(rng.choice([0.05, 0.3, 1.0])))
    else:                                           # a few huge triangles
        v = (rng.standard_normal((9, 3)) * 3).astype(np.float32)
        f = np.arange(9, dtype=np.int32).reshape(3, 3)
    c = rng.random(v.shape).astype(np.float32)
    rows, cols = [(480, 640), (120, 160), (97, 131), (33, 17)][int(rng.integers(0, 4))]
    cam = np.array([572.4 * cols / 640, 573.6 * rows / 480, cols * 0.51, rows * 0.49], np.float32)
    z_near = float(rng.choice([0.001, 0.05, 0.3]))
    poses = np.stack([_rand_pose(rng, float(rng.choice([-0.5, 0.0, 0.02, 0.2, 0.6, 1.5, 5.0]))) for _ in range(3)])
    kw = {}
    if rng.random() < 0.4:
        kw = dict(uv=(rng.standard_normal((len(v), 2)) * 2).astype(np.float32),
                  tex=rng.integers(0, 256, (int(rng.integers(1, 40)), int(rng.integers(1, 40)), 3)).astype(np.uint8))
    tag = f"render kind={kind} {rows}x{cols} near={z_near} tex={bool(kw)} nv={len(v)} nf={len(f)}"
    if kw:
        gc, gd = gpu.render_views_tex(v, c, kw["uv"], f, kw["tex"], cam, poses, rows=rows, cols=cols, z_near=z_near, z_far=4.0)
    else:
        gc, gd = gpu.render_views(v, c, f, cam, poses, rows=rows, cols=cols, z_near=z_near, z_far=4.0)
    for i, rt in enumerate(poses):
        oc, od = po.render(v, c, f, cam, rt, rows=rows, cols=cols, z_near=z_near, z_far=4.0, **kw)
        assert np.array_equal(gd[i], od), tag + f" view {i}: depth differs at {int((gd[i] != od).sum())} pixels"
        assert np.array_equal(gc[i], oc), tag + f" view {i}: colour differs at {int((gc[i] != oc).any(-1).sum())} pixels"
    return float((gd > 0).mean())


def run_patch(gpu, po, rng):
    """samplePatchWithScale + depth2normals at random centres / depths, incl. cuts clamped at the border and impossible ones"""
    from tests import scenes
    rows, cols = [(480, 640), (240, 320), (200, 333)][int(rng.integers(0, 3))]
    col, dep, _ = scenes.make_frame(int(rng.integers(1 << 20)), rows, cols)
    fx, fy = 572.4 * cols / 640, 573.6 * rows / 480
    n = 24
    ctr = np.stack([rng.integers(-20, cols + 20, n), rng.integers(-20, rows + 20, n)], 1).astype(np.int32)
    z = rng.choice([0.25, 0.4, 0.72, 1.0, 2.5, 0.0, -1.0, np.inf, np.nan], n).astype(np.float32)
    got, status = gpu.sample_patches(col[None], dep[None], ctr, z, fx, fy, frame_of=np.zeros(n, np.int32))
    nor = po.depth2normals(dep, fx, fy)
    for i in range(n):
        want = po.sample_patch(col, dep, nor, ctr[i, 0], ctr[i, 1], float(z[i]), fx, fy)
        tag = f"patch {rows}x{cols} centre {ctr[i].tolist()} z {z[i]}"
        if want is None:
            assert status[i] == 1 and not got[i].any(), tag + " should be rejected"
        else:
            assert status[i] == 0 and np.array_equal(got[i], want), tag + f" differs at {int((got[i] != want).sum())} values"
    return float((status == 0).mean())


def main(argv):
    n_cases = int(argv[1]) if len(argv) > 1 else 120
    seed = int(argv[2]) if len(argv) > 2 else 1
    from oracle import pyoracle as po
    po.build()
    import importlib
    gpu = importlib.import_module("3dposeestimation_b200.capi")
    synth = importlib.import_module("3dposeestimation_b200.synth")
    rng = np.random.default_rng(seed)
    bad = 0
    for path in (1, 0):
        gpu.set_option("knn_tensor", path)
        for c in knn_cases(rng, n_cases if path else n_cases // 3):
            try:
                run_knn(gpu, po, c)
            except Exception as e:  # noqa: BLE001
                bad += 1
                print("FAIL knn", "tensor" if path else "simt", c, "->", repr(e)[:300], flush=True)
        for _ in range(n_cases // 4):
            try:
                run_hamming(gpu, po, rng)
            except Exception as e:  # noqa: BLE001
                bad += 1
                print("FAIL hamming", "tensor" if path else "simt", repr(e)[:300], flush=True)
    gpu.set_option("knn_tensor", 1)
    for name in netdefs.VARIANTS:
        for _ in range(3):
            try:
                run_forward(gpu, po, synth, rng, name)
            except Exception as e:  # noqa: BLE001
                bad += 1
                print("FAIL forward", name, repr(e)[:300], flush=True)
    for _ in range(n_cases // 2):
        try:
            run_eval(gpu, po, synth, rng)
        except Exception as e:  # noqa: BLE001
            bad += 1
            print("FAIL", repr(e)[:400], flush=True)
    for fn in (run_render, run_patch):
        seen = []
        for _ in range(n_cases // 4):
            try:
                seen.append(fn(gpu, po, rng))
            except Exception as e:  # noqa: BLE001
                bad += 1
                print("FAIL", repr(e)[:400], flush=True)
        what = "covered pixels" if fn is run_render else "accepted cuts"
        print(f"{fn.__name__}: {len(seen)} cases ok, {what}: mean {np.mean(seen):.3f}, cases with none {int(np.sum(np.array(seen) == 0))}")
    print(f"fuzz: {bad} failing cases (seed {seed}, {n_cases} knn cases per path)")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
