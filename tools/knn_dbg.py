import importlib, os, sys
import numpy as np
sys.path.insert(0, "/root/repo")
capi = importlib.import_module("3dposeestimation_b200.capi")
nt, nq = 1000200, 4096
rng = np.random.default_rng(1)
t = rng.standard_normal((nt, 16)).astype(np.float32)
q = (t[rng.integers(0, nt, nq)] + 0.05 * rng.standard_normal((nq, 16))).astype(np.float32)
db = capi.TemplateDB(desc=t)
for k in (1, 2, 5):
    capi.set_option("knn_tensor", 0); i0, d0 = db.knn(q, k)
    capi.set_option("knn_tensor", 1)
    for rep in range(3):
        i1, d1 = db.knn(q, k)
        bad = np.argwhere(i0 != i1)
        print("k", k, "rep", rep, "mismatching entries", len(bad), "queries", len(set(bad[:, 0].tolist())), "ranks", np.bincount(bad[:, 1], minlength=k).tolist() if len(bad) else [])
        if len(bad):
            qi = bad[0, 0]
            print("  q", qi, "simt", i0[qi], d0[qi], "tensor", i1[qi], d1[qi])
