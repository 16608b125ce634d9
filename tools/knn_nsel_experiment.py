#!/usr/bin/env python
"""How many candidate blocks should the select pass re-rank?  Emulates the 8-GPU shard shape of bench.py on one GPU
(125,025 rows x 16, 32,768 queries) and the 1-GPU shape (1,000,200 rows, 4,096 queries), for k = 1 and 5; prints scan /
select times and the number of queries the certificate sends to the exact re-scan.  Run once per PE_KNN_NSEL_EXTRA value."""
import importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
capi = importlib.import_module("3dposeestimation_b200.capi")
net = capi.Net(named="manifold_rgbdnor_16"); net.fill(20240607, 1)
out = {"extra": os.environ.get("PE_KNN_NSEL_EXTRA", "3")}
for T, Q in ((125025, 32768), (1000200, 4096)):
    d_p = capi.DeviceArray((4096, 64, 64, 7)); d_db = capi.DeviceArray((T, 16))
    for first in range(0, T, 4096):       # descriptors of synthetic patches: the distribution the bench matches against
        m = min(4096, T - first)
        net.synth_patches_dev(20240607, first, m, 3334, d_p); net.compute_descriptors_dev(d_p, m, d_db, out_offset=first)
    d_q = capi.DeviceArray((Q, 16))
    for first in range(0, Q, 4096):
        net.synth_patches_dev(20240608, first, 4096, 3334, d_p); net.compute_descriptors_dev(d_p, 4096, d_q, out_offset=first)
    net.sync()
    db = capi.TemplateDB(d_desc=d_db, t_local=T, dim=16)
    for k in (1, 5):
        d_i, d_d = capi.DeviceArray((Q, k), np.int32), capi.DeviceArray((Q, k))
        for _ in range(3): db.knn_dev(d_q, Q, k, d_i, d_d)
        db.sync(); db.profile(True)
        for _ in range(10): db.knn_dev(d_q, Q, k, d_i, d_d)
        ms, calls = db.profile_read(); db.profile(False)
        out[f"T{T}_Q{Q}_k{k}"] = {"scan_ms": round(ms[0] / calls, 4), "select_ms": round(ms[1] / calls, 4), "flagged": db.knn_flagged()}
print(json.dumps(out))
