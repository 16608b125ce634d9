"""examples/evaluate: the evaluation half of the reference's program (datasetManager(config) -> generateDatasets ->
networkEvaluator::saveLog / saveConfusionMatrix) as a stand-alone C++ binary on the mirror headers.  Driven here the way a user
would: sample files + config.ini + prototxt + weights on disk in, log files out; the histogram row must equal what the C-ABI
chain gives for the sets the mirror's generateDatasets builds (recomputed here in numpy: split by closest template, crop)."""
import os
import subprocess

import numpy as np
import pytest

from tests import netdefs

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BINS = np.array([-1, 0, 10, 20, 40, 180], np.float32)
NAME = "manifold_rgbdnor_16"
MODELS = ["ape", "bowl", "cam"]


def _ang(a, b):
    d = np.minimum(1.0, np.abs((a * b).sum(-1)))
    return 2 * np.arccos(d.astype(np.float32))


def test_evaluate_binary_end_to_end(gpu, po, synth, tmp_path):
    exe = os.path.join(ROOT, "examples", "evaluate")
    assert os.path.exists(exe), "examples/evaluate is built by __graft_entry__.build()"
    n_t, n_real, n_synth = 30, 26, 12
    (tmp_path / "net").mkdir()
    (tmp_path / "h5").mkdir()
    (tmp_path / "net" / f"{NAME}.prototxt").write_text(netdefs.prototxt(NAME))
    ref = po.parse_prototxt(netdefs.prototxt(NAME))
    net = gpu.Net(text=netdefs.prototxt(NAME))
    net.set_blobs(synth.make_blobs(ref.blob_names, ref.blob_shapes, 3, mode=1))
    net.save_weights(str(tmp_path / "w.pew"))
    sets = {}
    for o, m in enumerate(MODELS):
        tl = synth.make_labels(50 + o, 1, n_t)
        tl[:, 0] = o
        tp = synth.make_patches(60 + o, o * n_t, n_t, 7, n_t)
        src = np.random.default_rng(o).integers(0, n_t, n_real)
        rl = synth.perturb_labels(tl[src], 70 + o, max_deg=12.0)
        rp = (tp[src] + 0.03 * np.random.default_rng(80 + o).standard_normal(tp[src].shape)).astype(np.float32)
        sl = synth.perturb_labels(tl[:n_synth], 90 + o, max_deg=5.0)
        sp = tp[:n_synth].copy()
        gpu.samples_write(str(tmp_path / "h5" / f"templates_{m}.pess"), tp, tl)
        gpu.samples_write(str(tmp_path / "h5" / f"realSamples_{m}.pess"), rp, rl)
        gpu.samples_write(str(tmp_path / "h5" / f"synthSamples_{m}.pess"), sp, sl)
        sets[m] = (tp, tl, rp, rl, sp, sl)
    cfg = tmp_path / "config.ini"
    cfg.write_text(f"""[paths]
dataset_path = {tmp_path}/
hdf5_path = {tmp_path}/h5/
network_path = {tmp_path}/net/
background_path = {tmp_path}/
output_path = {tmp_path}/

[input]
dataset_name = LineMOD
models = ape,bowl,cam
used_models = ape,bowl,cam
rotInv = 0,1,0
inplane = false
use_real = true
random_background = 0

[train]
net_name = {NAME}

[output]
kNN = 3
log_name = _ex.log
""")
    r = subprocess.run([exe, str(cfg), str(tmp_path / "w.pew"), "120"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    # the sets generateDatasets builds (src/datasetmanager.cpp:618-668): a real sample hangs on its closest template (first
    # minimum); of each template's samples the first ceil(n / 2) go to training, the rest to the test set; crop to the minimum
    test_sets = []
    for m in MODELS:
        tp, tl, rp, rl, sp, sl = sets[m]
        closest = np.array([int(np.argmin(_ang(rl[i, 1:5][None], tl[:, 1:5]))) for i in range(n_real)])
        test_idx = []
        for t in range(n_t):
            of = np.nonzero(closest == t)[0]
            test_idx += list(of[int(np.ceil(len(of) / 2.0)):])
        test_sets.append(test_idx)
    n_test = min(len(x) for x in test_sets)
    assert n_test >= 3
    head = r.stdout.splitlines()[0]
    assert head.startswith(f"objects 3  templates/object {n_t}") and head.endswith(f"test/object {n_test}"), head
    # histogram over the cropped test sets (the order inside a set does not matter for the histogram) through the C ABI
    tdesc = np.concatenate([net.compute_descriptors(sets[m][0]) for m in MODELS])
    tlab = np.concatenate([sets[m][1] for m in MODELS])
    qp = np.concatenate([sets[m][2][test_sets[o][:n_test]] for o, m in enumerate(MODELS)])
    ql = np.concatenate([sets[m][3][test_sets[o][:n_test]] for o, m in enumerate(MODELS)])
    db = gpu.TemplateDB(desc=tdesc, labels_all=tlab, poses_per_object=n_t)
    idx, _ = db.knn(net.compute_descriptors(qp), 3)
    want = db.histogram(idx, ql, np.array([0, 1, 0], np.int32), BINS)
    assert np.array_equal(want[:6], po.histogram(idx, ql, tlab, n_t, np.array([0, 1, 0], np.int32), BINS)[:6])
    log = (tmp_path / "log_ex.log").read_text().splitlines()
    assert log[0].split("\t")[:4] == ["mod", "iter", "epoch", "time"] and [l.split("\t")[0] for l in log[1:]] == ["test", "train"]
    row = log[1].split("\t")
    assert row[1] == "120"
    got_bins = np.array([float(x) for x in row[4:10]])
    np.testing.assert_allclose(got_bins, want[:6] * 100, rtol=0, atol=0.06)          # the log keeps 3 significant digits
    np.testing.assert_allclose([float(row[10]), float(row[11])], want[6:], rtol=2e-3, atol=1e-2)
    cm = (tmp_path / "cm_ex.log").read_text().splitlines()
    assert cm[0].split("\t") == MODELS and len(cm) == 4
    # a missing sample file is an error with the file's name, not a crash
    os.remove(tmp_path / "h5" / "templates_bowl.pess")
    r = subprocess.run([exe, str(cfg)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 1 and "templates_bowl" in r.stderr
