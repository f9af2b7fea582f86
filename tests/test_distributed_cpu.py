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


def _worker_owner_reduce(rank, world, port, iters, out_dir):
    """The data flow of the rank-mode exchange (csrc/mf_iter.cu, k_xchg_ll) on CPU: every rank
    sends the gradient rows of its shard to the rank that owns the item (blocks of
    slots_per_rank touched items, host/mf_prep.c:mf_exchange_layout); the owner adds the
    contributions in rank order -- ranks without a nonzero in the item count as +0.0 and send
    nothing --, applies, and sends its rows of R_new to every replica."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from recommender_system_b200 import api, synth
    inst = synth.random_instance(40, 37, 5, 0.15, iters, 0.002, seed=9)   # sparse: many items miss a rank
    U, I, F = inst.users, inst.items, inst.features
    cut = api.split_users(U, inst.entries, world)
    slot, contrib, per = api.exchange_layout(U, I, inst.entries, world)
    touched = np.flatnonzero(slot >= 0)                     # item of slot t (slots are in item order)
    assert np.array_equal(slot[touched], np.arange(len(touched)))
    lo, hi = int(cut[rank]), int(cut[rank + 1])
    L, R = api.random_fill_LR(U, I, F)
    mine_lo, mine_hi = rank * per, min(len(touched), (rank + 1) * per)
    for _ in range(iters):
        G = _shard_iteration(inst.entries, lo, hi, inst.alpha, L, R, I, F)
        # "push": the rows of every owner's block that this rank contributes to
        send = []
        for o in range(world):
            rows = {}
            for t in range(o * per, min(len(touched), (o + 1) * per)):
                if (contrib[t] >> rank) & 1:
                    rows[t] = G[touched[t]].copy()
            send.append(rows)
        # (gloo has no object all-to-all: gather everybody's parcels and keep this rank's)
        parcels = [None] * world
        dist.all_gather_object(parcels, send)
        recv = [parcels[h][rank] for h in range(world)]     # what rank h sent to this rank
        # owner reduce, rank order; non-contributors add +0.0
        new_rows = {}
        for t in range(mine_lo, mine_hi):
            s = recv[0][t] if (contrib[t] & 1) else np.zeros(F)
            for h in range(1, world):
                s = s + (recv[h][t] if (contrib[t] >> h) & 1 else np.zeros(F))
            assert ((contrib[t] >> np.arange(world)) & 1).tolist() == [t in recv[h] for h in range(world)]
            new_rows[t] = R[touched[t]] + s
        # broadcast of the own block into every replica
        blocks = [None] * world
        dist.all_gather_object(blocks, new_rows)
        R = R.copy()
        for blk in blocks:
            for t, row in blk.items():
                R[touched[t]] = row
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, L[lo:hi].copy()))
    for (a, b, rows) in parts:
        L[a:b] = rows
    np.save(os.path.join(out_dir, "L%d.npy" % rank), L)
    np.save(os.path.join(out_dir, "R%d.npy" % rank), R)
    dist.destroy_process_group()


def test_two_rank_owner_reduce_matches_sharded_oracle(built_lib, oracle, tmp_path):
    """World size 2 on gloo: the owner-reduces / owner-broadcasts exchange with the library's own
    layout gives the oracle's N-shard doubles bit for bit, replicas identical."""
    from recommender_system_b200 import synth
    world, iters = 2, 6
    mp.spawn(_worker_owner_reduce, args=(world, _free_port(), iters, str(tmp_path)), nprocs=world, join=True)
    inst = synth.random_instance(40, 37, 5, 0.15, iters, 0.002, seed=9)
    cut = built_lib.split_users(inst.users, inst.entries, world)
    L0, R0 = oracle.init_LR(inst.users, inst.items, inst.features)
    oracle.factorize_sharded(inst.users, inst.items, inst.features, inst.entries, iters, inst.alpha,
                             L0, R0, cut)
    Ls = [np.load(tmp_path / ("L%d.npy" % r)) for r in range(world)]
    Rs = [np.load(tmp_path / ("R%d.npy" % r)) for r in range(world)]
    assert np.array_equal(Rs[0], Rs[1]) and np.array_equal(Ls[0], Ls[1])
    assert np.array_equal(Rs[0].view(np.uint64), R0.view(np.uint64))
    assert np.array_equal(Ls[0].view(np.uint64), L0.view(np.uint64))


def test_exchange_layout_properties(built_lib):
    """mf_exchange_layout: every touched item has a slot in item order, blocks of slots_per_rank
    cover them, and contrib is exactly the set of ranks whose user range holds a nonzero."""
    from recommender_system_b200 import synth
    for (U, I, dens, n) in ((50, 30, 0.1, 3), (7, 90, 0.05, 8), (200, 11, 0.5, 16), (5, 5, 0.0, 2)):
        inst = synth.random_instance(U, I, 3, dens, 1, 0.01, seed=U + n)
        cut = built_lib.split_users(U, inst.entries, n)
        slot, contrib, per = built_lib.exchange_layout(U, I, inst.entries, n)
        cols = np.unique(inst.entries["col"])
        assert np.array_equal(np.flatnonzero(slot >= 0), cols) and len(contrib) == len(cols)
        assert per == max(1, -(-len(cols) // n))
        want = np.zeros(len(cols), dtype=np.int64)
        for e in inst.entries:
            r = int(np.searchsorted(cut, e["row"], side="right")) - 1
            while cut[r] == cut[r + 1]:
                r += 1
            want[slot[e["col"]]] |= 1 << r
        assert np.array_equal(want, contrib)
