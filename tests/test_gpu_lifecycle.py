"""Handle lifecycle through the C ABI: create -> use -> destroy, many times, must give the device memory back (SURVEY 8b
"Ownership": the library owns device buffers behind handles; nothing may outlive pe_*_destroy).  Free memory is read with
torch.cuda.mem_get_info (plumbing only)."""
import gc

import numpy as np
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu
BINS = np.array([-1, 0, 10, 20, 40, 180], np.float32)


def _one_round(gpu, synth, i):
    net = gpu.Net(named="manifold_rgbdnor_16" if i % 2 == 0 else "manifold_rgbdnor_16_bin")
    net.fill(i, 1)
    n_obj, nr_t = 5, 400
    tl = synth.make_labels(i, n_obj, nr_t)
    x = synth.make_patches(i, 0, 96, 7, 10)
    d = net.compute_descriptors(x)
    t = np.random.default_rng(i).standard_normal((n_obj * nr_t, d.shape[1])).astype(np.float32)
    db = gpu.TemplateDB(desc=t, labels_all=tl, poses_per_object=nr_t)
    idx, _ = db.knn(d, 5)
    ql = tl[idx[:, 0]].copy()
    db.histogram(idx, ql, np.zeros(n_obj, np.int32), BINS)
    db.confusion(idx, ql, np.arange(n_obj, dtype=np.int32), n_obj, 96 // n_obj + 1)
    codes = gpu.TemplateDB(codes=np.random.default_rng(i).integers(0, 1 << 32, (3000, 4), dtype=np.uint64).astype(np.uint32))
    codes.knn_hamming(np.zeros((8, 4), np.uint32), 3)
    v, f = scenes.icosphere(2, 0.08)
    gpu.render_views(v, (v / 0.2 + 0.5).astype(np.float32), f, np.array([572.4, 573.6, 325.3, 242.0], np.float32),
                     gpu.sphere_views(0, (0.5, 1, 0.5), (0, 1, 0))[:2])
    da = gpu.DeviceArray((1 << 20,))
    del da
    for h in (codes, db, net):
        h.close()


def test_create_use_destroy_returns_device_memory(gpu, synth):
    torch = pytest.importorskip("torch")
    for i in range(3):                       # first rounds: context, module load, the allocator's own pools
        _one_round(gpu, synth, i)
    gc.collect()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    import psutil
    rss0 = psutil.Process().memory_info().rss
    for i in range(3, 28):
        _one_round(gpu, synth, i)
    gc.collect()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    lost = free0 - free1
    assert lost < 16 << 20, f"{lost / 2**20:.1f} MB of device memory not returned after 25 create / use / destroy rounds"
    grown = psutil.Process().memory_info().rss - rss0        # pinned staging buffers are host memory of this process
    assert grown < 256 << 20, f"host memory grew by {grown / 2**20:.0f} MB over 25 rounds"


def test_distinct_handles_from_distinct_threads(gpu, synth):
    """SURVEY 8b "Threading": a handle belongs to one thread at a time, distinct handles may run on distinct threads (one
    stream per handle).  Four threads, each with its own net + DB, run forward -> k-NN -> histogram concurrently (ctypes
    drops the GIL during the calls); every thread must get exactly what it gets alone."""
    import threading
    names = ["manifold_rgbdnor_16", "manifold_rgbd_16", "manifold_rgbdnor_16_bin", "manifold_rgb_16"]
    n_obj, nr_t = 4, 500

    def work(i, rounds):
        net = gpu.Net(named=names[i])
        net.fill(10 + i, 1)
        x = synth.make_patches(20 + i, 0, 200, 7, 10)
        tl = synth.make_labels(30 + i, n_obj, nr_t)
        out = []
        for r in range(rounds):
            d = net.compute_descriptors(x[r * 7:r * 7 + 150])
            t = np.random.default_rng(40 + i).standard_normal((n_obj * nr_t, d.shape[1])).astype(np.float32) * np.abs(d).mean()
            db = gpu.TemplateDB(desc=t, labels_all=tl, poses_per_object=nr_t)
            idx, dist = db.knn(d, 3)
            h = db.histogram(idx, tl[idx[:, 0]].copy(), np.zeros(n_obj, np.int32), BINS)
            db.close()
            out.append((d, idx, dist, h))
        net.close()
        return out

    alone = [work(i, 4) for i in range(4)]
    got = [None] * 4
    errs = []

    def run(i):
        try:
            got[i] = work(i, 4)
        except Exception as e:  # noqa: BLE001
            errs.append((i, repr(e)))

    threads = [threading.Thread(target=run, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(300)
    assert not errs, errs
    for i in range(4):
        for a, b in zip(alone[i], got[i]):
            for u, v in zip(a, b):
                assert np.array_equal(u, v, equal_nan=True), f"thread {i} got another result than alone"
