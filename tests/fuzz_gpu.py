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
    print(f"fuzz: {bad} failing cases (seed {seed}, {n_cases} knn cases per path)")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
