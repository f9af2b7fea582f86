"""The N>1 scheme on CPU: two gloo ranks, users sharded with the library's own
nonzero-balanced split, R replicated, partial R gradient all-reduced every
iteration, R_new = R_old + sum (what rank mode does with NCCL).  Each rank
restates the kernels' arithmetic in numpy for its shard; the result must be
bit-identical to the oracle's N-shard ordering and the replicas must agree."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, load_instance


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _shard_iteration(entries, lo, hi, alpha, L, R, items, F):
    """One iteration for the users [lo, hi): returns the shard's gradient G (from 0.0,
    file order) and updates the shard's rows of L in place -- mf_loop.cuh's order."""
    Ls = L.copy()
    G = np.zeros((items, F))
    for n in range(len(entries)):
        i, j, a = int(entries[n]["row"]), int(entries[n]["col"]), float(entries[n]["value"])
        if not (lo <= i < hi):
            continue
        s = 0.0
        for k in range(F):
            s = s + Ls[i, k] * R[j, k]
        e = (alpha * 2) * (a - s)
        for k in range(F):
            L[i, k] = L[i, k] + e * R[j, k]
            G[j, k] = G[j, k] + e * Ls[i, k]
    return G


def _worker(rank, world, port, iters, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from recommender_system_b200 import api
    from recommender_system_b200.io import read_instance
    inst = read_instance(os.path.join(ROOT, "tests", "golden", "inst30-40-10-2-10.in"))
    cut = api.split_users(inst.users, inst.entries, world)
    lo, hi = int(cut[rank]), int(cut[rank + 1])
    L, R = api.random_fill_LR(inst.users, inst.items, inst.features)
    for _ in range(iters):
        G = _shard_iteration(inst.entries, lo, hi, inst.alpha, L, R, inst.items, inst.features)
        t = torch.from_numpy(G)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        R = R + t.numpy()
    # L rows live on their owner; gather them on every rank
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, L[lo:hi].copy()))
    for (a, b, rows) in parts:
        L[a:b] = rows
    np.save(os.path.join(out_dir, "L%d.npy" % rank), L)
    np.save(os.path.join(out_dir, "R%d.npy" % rank), R)
    dist.destroy_process_group()


def test_two_rank_gradient_allreduce_matches_sharded_oracle(built_lib, oracle, tmp_path):
    world, iters = 2, 12
    mp.spawn(_worker, args=(world, _free_port(), iters, str(tmp_path)), nprocs=world, join=True)
    inst = load_instance("inst30-40-10-2-10")
    cut = built_lib.split_users(inst.users, inst.entries, world)
    L0, R0 = oracle.init_LR(inst.users, inst.items, inst.features)
    oracle.factorize_sharded(inst.users, inst.items, inst.features, inst.entries, iters, inst.alpha,
                             L0, R0, cut)
    Ls = [np.load(tmp_path / ("L%d.npy" % r)) for r in range(world)]
    Rs = [np.load(tmp_path / ("R%d.npy" % r)) for r in range(world)]
    assert np.array_equal(Rs[0], Rs[1]) and np.array_equal(Ls[0], Ls[1])   # replicas agree
    assert np.array_equal(Rs[0], R0) and np.array_equal(Ls[0], L0)         # == oracle ordering
