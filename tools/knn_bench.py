#!/usr/bin/env python
"""kNN-only micro-benchmark: 4096 queries vs a random 1,000,200 x 16 DB (device-resident), both scan paths.
   PE_TC_DEBUG=1 prints the tensor scan's per-role cycle counters."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
capi = importlib.import_module("3dposeestimation_b200.capi")
nt, nq, k = int(os.environ.get("NT", 1000200)), int(os.environ.get("NQ", 4096)), int(os.environ.get("K", 1))
rng = np.random.default_rng(1)
t = rng.standard_normal((nt, 16)).astype(np.float32)
q = (t[rng.integers(0, nt, nq)] + 0.05 * rng.standard_normal((nq, 16))).astype(np.float32)
db = capi.TemplateDB(desc=t)
d_q = capi.DeviceArray.from_host(q)
d_i = capi.DeviceArray((nq, k), np.int32); d_d = capi.DeviceArray((nq, k), np.float32)
res = {}
for path in (1, 0):
    capi.set_option("knn_tensor", path)
    for _ in range(2):
        db.knn_dev(d_q, nq, k, d_i, d_d)
    db.sync()
    tm = capi.Timer(0); tm.start(db)
    for _ in range(5):
        db.knn_dev(d_q, nq, k, d_i, d_d)
    tm.stop(db)
    ms = tm.elapsed_ms() / 5
    res[path] = d_i.to_host().copy()
    print(f"knn path={'tensor' if path else 'simt'} nq={nq} nt={nt} k={k}: {ms:.3f} ms/call, {nq / ms * 1e3:.0f} queries/s")
print("paths agree:", np.array_equal(res[1], res[0]))
