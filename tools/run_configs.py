#!/usr/bin/env python
"""Throughput of the BASELINE.json configurations other than the bench.py headline (configs[1]):

  python tools/run_configs.py --config 1            # manifold.prototxt (C=4, D=3), 15 x 1,300 templates, 9,000 queries, 1-NN
  python tools/run_configs.py --config 4            # manifold_rgbdnor_16_bin: forward + sign bits + Hamming k-NN vs 1,000,200 codes
  python tools/run_configs.py --config 5            # rgbdnor_16, 1,000,200 templates, 30,000 queries, k = 5 + angular-error histogram
  python tools/run_configs.py --config 6            # rgbdnor_16_bin float descriptors (D = 128), L2 k-NN vs 1,000,200 rows (tensor scan)
  torchrun --nproc-per-node N tools/run_configs.py --config 3   # 125 x 10,000 templates row-sharded over N GPUs (also 4, 5, 6)

Inputs are the seeded synthetic patches of the bench (device generator), resident in HBM; times are CUDA events on
the launching stream, max over ranks; one JSON line per configuration on rank 0.  Parity for the same configurations
is the job of tests/test_gpu_parity.py - this tool only measures."""
import argparse
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SEED = 20240607
BINS = np.array([-1, 0, 10, 20, 40, 180], np.float32)

CONFIGS = {
    1: dict(net="manifold", c=4, d=3, n_obj=15, poses=1300, q_poses=600, batch=9000, k=1, hamming=False, hist=True),
    3: dict(net="manifold_rgbdnor_16", c=7, d=16, n_obj=125, poses=10000, q_poses=None, batch=4096, k=1, hamming=False, hist=False),
    4: dict(net="manifold_rgbdnor_16_bin", c=7, d=128, n_obj=300, poses=3334, q_poses=None, batch=4096, k=5, hamming=True, hist=False),
    # Q = 30,000 (300 x 100) queries in total, split over the ranks (the same query set at every N, so the
    # histogram of an N-GPU run must equal the single-GPU one), rotInv 80 / 10 / 10 %
    5: dict(net="manifold_rgbdnor_16", c=7, d=16, n_obj=300, poses=3334, q_poses=100, batch=30000, k=5, hamming=False, hist=True, split=True),
    # the float bin_ip2 descriptors the reference's evaluator L2-matches (networkevaluator.cpp:195,265): D = 128
    6: dict(net="manifold_rgbdnor_16_bin", c=7, d=128, n_obj=300, poses=3334, q_poses=None, batch=4096, k=1, hamming=False, hist=False),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True, choices=sorted(CONFIGS))
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    a = ap.parse_args()
    cfg = CONFIGS[a.config]
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(dev)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{dev}"))
    capi = importlib.import_module("3dposeestimation_b200.capi")
    synth = importlib.import_module("3dposeestimation_b200.synth")
    shard_rows = importlib.import_module("3dposeestimation_b200.sharding").shard_rows
    net = capi.Net(named=cfg["net"], device=dev)
    net.fill(SEED, 1)
    C, D, B, K = cfg["c"], cfg["d"], cfg["batch"], cfg["k"]
    if cfg.get("split"):                                    # a fixed query set split over the ranks
        assert B % world == 0
        B //= world
    T = cfg["n_obj"] * cfg["poses"]
    lo, hi = shard_rows(rank, world, T)
    chunk = 4096
    d_in = capi.DeviceArray((max(B, chunk), 64, 64, C), device=dev)
    d_db = capi.DeviceArray((hi - lo, D), device=dev)
    for first in range(lo, hi, chunk):                      # untimed: template descriptors of this rank's rows
        m = min(chunk, hi - first)
        net.synth_patches_dev(SEED, first, m, cfg["poses"], d_in)
        net.compute_descriptors_dev(d_in, m, d_db, out_offset=first - lo)
    net.sync()
    labels = capi.synth_labels(SEED, 0, T, cfg["poses"])
    comm = None
    if world > 1:
        uid = [capi.Comm.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        comm = capi.Comm(rank, world, uid[0], dev)
    if cfg["hamming"]:
        words = D // 32
        d_codes = capi.DeviceArray((hi - lo, words), np.uint32, device=dev)
        net.binarize_dev(d_db, hi - lo, D, d_codes)
        net.sync()
        db = capi.TemplateDB(codes=d_codes, labels_all=labels, poses_per_object=cfg["poses"], row_offset=lo, t_total=T, device=dev)
    else:
        db = capi.TemplateDB(d_desc=d_db, t_local=hi - lo, dim=D, row_offset=lo, t_total=T, labels_all=labels,
                             poses_per_object=cfg["poses"], device=dev)
    db.share_stream(net)
    qp = cfg["q_poses"] or cfg["poses"]
    q_labels = None
    if cfg["hist"]:
        # evaluation queries (configs 1 and 5): query g of class c = g // qp is the patch of template pose g % qp of that class
        # (so its nearest template is known: random-weight descriptors of *other* patches of a class do not land in that
        # class), its pose label that template's pose rotated by up to 45 degrees: the histogram then spreads over all bins.
        # The whole query set is defined by global ids, so it is identical at every N.
        assert B % qp == 0 or qp % B == 0 or True
        g0 = rank * B
        g = g0
        while g < g0 + B:
            c, j = divmod(g, qp)
            run = min(qp - j, g0 + B - g)
            net.synth_patches_dev(SEED, c * cfg["poses"] + j, run, cfg["poses"], d_in, out_offset=g - g0)
            g += run
        ids = np.arange(B * world)
        src = (ids // qp) * cfg["poses"] + (ids % qp)
        q_labels = np.ascontiguousarray(synth.perturb_labels(labels[src], SEED + 1, max_deg=45.0)[rank * B:(rank + 1) * B])
    else:
        net.synth_patches_dev(SEED + 1, rank * B, B, qp, d_in)                   # this rank's query patches
    nq_all = B * world
    d_desc = capi.DeviceArray((B, D), device=dev)
    d_all = capi.DeviceArray((nq_all, D), device=dev) if world > 1 else d_desc
    d_idx = capi.DeviceArray((nq_all, K), np.int32, device=dev)
    d_dist = capi.DeviceArray((nq_all, K), np.int32 if cfg["hamming"] else np.float32, device=dev)
    d_bits = capi.DeviceArray((nq_all, max(1, D // 32)), np.uint32, device=dev) if cfg["hamming"] else None

    def forward():
        net.compute_descriptors_dev(d_in, B, d_desc)
        if world > 1:
            comm.allgather_f32_dev(net, d_desc, B * D, d_all)

    def match():
        if cfg["hamming"]:
            net.binarize_dev(d_all, nq_all, D, d_bits)
            db.knn_hamming_sharded_dev(comm, d_bits, nq_all, K, d_idx, d_dist)
        else:
            db.knn_sharded_dev(comm, d_all, nq_all, K, d_idx, d_dist)

    def timed(fn):
        for _ in range(a.warmup):
            fn()
        net.sync()
        if world > 1:
            dist.barrier()
        t = capi.Timer(dev)
        t.start(net)
        for _ in range(a.steps):
            fn()
        t.stop(net)
        ms = t.elapsed_ms() / a.steps
        if world > 1:
            import torch
            tt = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{dev}")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        return ms

    ms_f = timed(forward)
    ms_m = timed(match)
    ms_both = timed(lambda: (forward(), match()))
    out = {"config": a.config, "net": cfg["net"], "n_gpus": world, "templates": T, "batch_per_gpu": B, "k": K,
           "matcher": "hamming" if cfg["hamming"] else "l2", "descriptors_per_s": B * world / (ms_f * 1e-3),
           "knn_queries_per_s": nq_all / (ms_m * 1e-3), "patches_per_s_forward_plus_knn": B * world / (ms_both * 1e-3),
           "ms_forward": ms_f, "ms_knn": ms_m, "ms_step": ms_both,
           "knn_flagged": None if cfg["hamming"] else int(db.knn_flagged()),     # queries of the last call sent to the exact re-scan
           "paths": {"forward": "tcgen05" if capi.get_option("forward_tensor") else "simt",
                     "knn": ("tcgen05 +-1 e4m3" if capi.get_option("knn_tensor") and D <= 128 and K <= 5 else "simt xor+popc") if cfg["hamming"]
                     else ("tcgen05" if capi.get_option("knn_tensor") and D <= 128 and K <= 16 else "simt")}}
    if cfg["hist"]:
        # computeHistogram's body over the whole query set: every rank reduces its own slice of the merged neighbours,
        # bins / sum all-reduced, per-query errors all-gathered for the median (pe_histogram_sharded_dev)
        rot_inv = synth.rot_inv_pattern(cfg["n_obj"])                    # 80 % / 10 % / 10 % as config.ini:10
        d_ql = capi.DeviceArray.from_host(q_labels, device=dev)
        d_own = capi.DeviceArray((B, K), np.int32, device=dev)
        own_idx = np.ascontiguousarray(d_idx.to_host()[rank * B:(rank + 1) * B])      # kept alive across the C call
        capi.check(capi.lib().pe_memcpy_h2d(dev, d_own.ptr, own_idx.ctypes.data, own_idx.nbytes))
        net.sync()
        h = db.histogram_sharded_dev(comm, d_own, d_ql, B, K, rot_inv, BINS)       # first call: sizes the scratch (cudaMalloc)
        t = capi.Timer(dev)
        t.start(db)
        for _ in range(5):
            h = db.histogram_sharded_dev(comm, d_own, d_ql, B, K, rot_inv, BINS)
        t.stop(db)
        out["ms_histogram"] = t.elapsed_ms() / 5
        out["histogram"] = [float(x) for x in h]
        out["histogram_queries"] = B * world
        out["rot_inv"] = "0 / 1 / 2 for 80 / 10 / 10 % of the classes (synth.rot_inv_pattern)"
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
