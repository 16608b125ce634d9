"""N > 1 path on CPU: two gloo ranks (127.0.0.1) run the sharded-matcher PROTOCOL - row ranges, global
indices, all-gather of per-shard candidates, merge rule - with the oracle standing in for the CUDA scan
(the oracle is the checker here, never the product).  The CUDA kernels of the same protocol are covered
by the GPU tests (test_sharded_scan_plus_merge_equals_unsharded, test_full_size_db_properties)."""
import importlib
import os
import socket

import numpy as np
import pytest

sharding = importlib.import_module("3dposeestimation_b200.sharding")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, nt, nq, k, dim, out_dir):
    import torch
    import torch.distributed as dist
    from oracle import pyoracle as po
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(7)                      # same data on every rank
    t = rng.standard_normal((nt, dim)).astype(np.float32)
    t[nt - 1] = t[0]                                    # a duplicate pair that straddles the two shards
    q = (t[rng.integers(0, nt, nq)] + 0.1 * rng.standard_normal((nq, dim))).astype(np.float32)
    q[0] = t[0]
    lo, hi = sharding.shard_rows(rank, world, nt)
    li, ld = po.knn_l2(q, t[lo:hi], k)                  # local scan of this rank's rows
    gi = np.where(li >= 0, li + lo, -1).astype(np.int32)
    all_i = [torch.empty((nq, k), dtype=torch.int32) for _ in range(world)]
    all_d = [torch.empty((nq, k), dtype=torch.float32) for _ in range(world)]
    dist.all_gather(all_i, torch.from_numpy(gi))
    dist.all_gather(all_d, torch.from_numpy(np.ascontiguousarray(ld)))
    mi, md = sharding.merge_rule([x.numpy() for x in all_i], [x.numpy() for x in all_d], k)
    oi, od = po.knn_l2(q, t, k)
    ok = bool(np.array_equal(mi, oi) and np.array_equal(md, od) and mi[0, 0] == 0 and (k < 2 or mi[0, 1] == nt - 1))
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        with open(os.path.join(out_dir, "result.txt"), "w") as f:
            f.write(str(int(flag.item())))
    dist.destroy_process_group()


def test_shard_rows_cover_and_balance():
    for total in (0, 1, 7, 1000200, 1250000):
        for world in (1, 2, 4, 8):
            spans = [sharding.shard_rows(r, world, total) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_rows(2, 2, 10)


@pytest.mark.parametrize("k", [1, 3])
def test_two_gloo_ranks_match_unsharded_oracle(po, tmp_path, k):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 1501, 64, k, 16, str(tmp_path)), nprocs=2, join=True)
    assert open(tmp_path / "result.txt").read() == "1"


def _hist_worker(rank, world, port, out_dir):
    """the sharded-evaluation protocol of pe_histogram_sharded_dev on CPU: every rank reduces its slice of the queries,
    bins / sums are combined, per-query errors gathered for the median == one unsharded pass"""
    import torch.distributed as dist
    from oracle import pyoracle as po
    synth = importlib.import_module("3dposeestimation_b200.synth")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_obj, nr_t, nr_q, k = 6, 50, 20, 3
    tl = synth.make_labels(5, n_obj, nr_t)
    src = (np.arange(nr_q) * 3) % nr_t
    ql = synth.perturb_labels(tl.reshape(n_obj, nr_t, 8)[:, src].reshape(-1, 8), 6, max_deg=30.0)
    rng = np.random.default_rng(3)
    nq = n_obj * nr_q
    own = (np.repeat(np.arange(n_obj), nr_q) * nr_t)[:, None] + rng.integers(0, nr_t, (nq, k))
    idx = np.where(rng.random((nq, k)) < 0.6, own, rng.integers(0, n_obj * nr_t, (nq, k))).astype(np.int32)
    rot = synth.rot_inv_pattern(n_obj)
    bins = np.array([-1, 0, 10, 20, 40, 180], np.float32)
    lo, hi = rank * nq // world, (rank + 1) * nq // world
    h_r, e_r = po.histogram(idx[lo:hi], ql[lo:hi], tl, nr_t, rot, bins, want_err=True)
    gathered = [None] * world
    dist.all_gather_object(gathered, (h_r, e_r))
    merged = sharding.merge_histograms([g[0] for g in gathered], [g[1] for g in gathered], bins)
    want = po.histogram(idx, ql, tl, nr_t, rot, bins)
    ok = bool(np.array_equal(merged[:6], want[:6]) and np.allclose(merged[6:], want[6:], atol=1e-4))
    if rank == 0:
        with open(os.path.join(out_dir, "hist.txt"), "w") as f:
            f.write("1" if ok else f"0 {merged} {want}")
    dist.destroy_process_group()


def test_two_gloo_ranks_sharded_histogram_equals_unsharded(po, tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_hist_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert open(tmp_path / "hist.txt").read() == "1"
