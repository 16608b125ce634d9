#!/usr/bin/env python
"""Multi-GPU check of the row-sharded matcher (run under torchrun, one rank per GPU):
   peer-window all-gather == NCCL all-gather == unsharded scan, for a few shapes, many calls in a row (epoch / parity
   reuse), plus timing of both all-gather kinds.  Prints one line on rank 0 and exits non-zero on a mismatch.

   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/p2p_check.py"""
import importlib, os, sys, time
import numpy as np
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
capi = importlib.import_module("3dposeestimation_b200.capi")
sharding = importlib.import_module("3dposeestimation_b200.sharding")

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
dist.init_process_group("gloo")
uid = [capi.Comm.unique_id() if rank == 0 else None]
dist.broadcast_object_list(uid, src=0)
comm = capi.Comm(rank, world, uid[0], local)
rng = np.random.default_rng(7)
T, D = 200_003, 16
t = rng.standard_normal((T, D)).astype(np.float32)
lo, hi = sharding.shard_rows(rank, world, T)
shard = capi.TemplateDB(desc=t[lo:hi], row_offset=lo, t_total=T, device=local)
full = capi.TemplateDB(desc=t, device=local) if rank == 0 else None
ok = True
capi.set_option("p2p", 1)
active = comm.p2p_active()          # windows mapped on every rank (with the option on)
for nq, k in ((1, 1), (257, 3), (4096, 1), (8192, 5), (4099, 5)):
    q = (t[rng.integers(0, T, nq)] + 0.05 * rng.standard_normal((nq, D))).astype(np.float32)
    res = {}
    for mode in (1, 0, 1, 1, 0):                       # alternate the two kinds, repeat (window parity reuse)
        capi.set_option("p2p", mode)
        res.setdefault(mode, []).append(shard.knn_sharded(comm, q, k))
    base_i, base_d = res[1][0]
    for m in res:
        for i, d in res[m]:
            ok &= bool(np.array_equal(i, base_i) and np.array_equal(d, base_d))
    if rank == 0:
        fi, fd = full.knn(q, k)
        ok &= bool(np.array_equal(fi, base_i) and np.allclose(fd, base_d, rtol=2e-6))
comm.p2p_active()                                       # raises if an all-gather timed out
# timing: all-gather of 4096 x 16 floats (the descriptor exchange of bench.py) on the device
net = capi.Net(named="manifold_rgbdnor_16", device=local)
net.fill(1, 1)
d_s = capi.DeviceArray((4096, 16), device=local); d_r = capi.DeviceArray((4096 * world, 16), device=local)
times = {}
for mode in (1, 0):
    capi.set_option("p2p", mode)
    for _ in range(5):
        comm.allgather_f32_dev(net, d_s, 4096 * 16, d_r)
    net.sync(); dist.barrier()
    tm = capi.Timer(local); tm.start(net)
    for _ in range(50):
        comm.allgather_f32_dev(net, d_s, 4096 * 16, d_r)
    tm.stop(net)
    times[mode] = tm.elapsed_ms() / 50 * 1e3
flags = [None] * world
dist.all_gather_object(flags, ok)
if rank == 0:
    print(f"p2p_check world={world} peer_windows={active} all_equal={all(flags)} allgather 262 KB/rank: peer kernel {times[1]:.1f} us, NCCL {times[0]:.1f} us")
capi.set_option("p2p", 0)
comm.close()
dist.destroy_process_group()
sys.exit(0 if all(flags) else 1)
