"""Worker of tests/test_gpu_parity.py::test_rank_mode_*: one process per GPU
(launched with torch.distributed.run), users sharded by rank, R gradient
all-reduced by NCCL inside the C-ABI library.  Rank 0 writes the result."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recommender_system_b200 import api, synth  # noqa: E402


def main():
    out_path = sys.argv[1]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    dist.init_process_group("gloo")
    uid = [api.Comm.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    comm = api.Comm(world, rank, uid[0], local)
    how = sys.argv[2] if len(sys.argv) > 2 else "ll"
    if how == "timeout":
        return timeout_case(out_path, rank, world, comm)
    if how == "soak":
        return soak_case(out_path, rank, world, comm, int(sys.argv[3]), int(sys.argv[4]))

    def set_how(p):
        if how != "ll":
            p.set_option("exchange", {"nccl": 0, "pull": 2, "push": 3}[how])

    # a smaller problem first: the communicator keeps its exchange buffers, the plans below
    # find them too small (re-created) and then, the second time round, reuse them as they are
    small = synth.random_instance(60, 50, 6, 0.2, 5, 0.002, seed=5)
    Ls, Rs = api.random_fill_LR(small.users, small.items, small.features)
    with api.Plan(small.users, small.items, small.features, small.entries, small.alpha, comm=comm) as p0:
        set_how(p0)
        p0.set_factors(Ls, Rs)
        p0.iterate(5)
        p0.sync()
    inst = synth.random_instance(300, 200, 20, 0.1, 25, 0.002, seed=33)
    L, R = api.random_fill_LR(inst.users, inst.items, inst.features)
    L1, R1 = L.copy(), R.copy()
    with api.Plan(inst.users, inst.items, inst.features, inst.entries, inst.alpha, comm=comm) as p1:
        set_how(p1)
        p1.set_factors(L1, R1)
        p1.iterate(25)
        p1.get_factors(L1, R1)
    plan = api.Plan(inst.users, inst.items, inst.features, inst.entries, inst.alpha, comm=comm)
    set_how(plan)
    kind = plan.exchange_kind()
    plan.set_factors(L, R)
    plan.iterate(17)           # graph path (16) + one direct iteration
    plan.iterate(8)
    lo, hi = plan.user_range(0)
    plan.get_factors(L, R)
    best = plan.recommend()
    plan.close()
    # the plan that reused the buffers of the one before it gives the same doubles
    assert np.array_equal(R.view(np.uint64), R1.view(np.uint64)), "second plan on the same communicator differs"
    assert np.array_equal(L[lo:hi].view(np.uint64), L1[lo:hi].view(np.uint64))
    comm.close()
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, L[lo:hi].copy(), best[lo:hi].copy(), R.copy()))
    if rank == 0:
        best_all = np.full(inst.users, -2, dtype=np.int32)
        for (a, b, rows, bb, _) in parts:
            L[a:b] = rows
            best_all[a:b] = bb
        replicas_equal = all(np.array_equal(parts[0][4].view(np.uint64), p[4].view(np.uint64)) for p in parts)
        np.savez(out_path, L=L, R=R, best=best_all, replicas_equal=replicas_equal, exchange_kind=kind,
                 cut=np.array([p[0] for p in parts] + [inst.users]))
    dist.destroy_process_group()


def timeout_case(out_path, rank, world, comm):
    """Rank 0 runs two iterations more than the others: its exchange waits for contributions
    that never come; after comm_timeout_ms the wait gives up and the sync reports it."""
    inst = synth.random_instance(120, 80, 8, 0.2, 4, 0.002, seed=7)
    L, R = api.random_fill_LR(inst.users, inst.items, inst.features)
    code, msg = 0, ""
    plan = api.Plan(inst.users, inst.items, inst.features, inst.entries, inst.alpha, comm=comm)
    plan.set_option("graph", 0)
    plan.set_option("comm_timeout_ms", 300)
    plan.set_factors(L, R)
    plan.iterate(2)
    plan.sync()
    if rank == 0:
        plan.iterate(2)
        try:
            plan.sync()
        except api.MatFactError as e:
            code, msg = e.code, str(e)
    dist.barrier()
    plan.close()
    comm.close()
    if rank == 0:
        np.savez(out_path, code=code, message=msg)
    dist.destroy_process_group()


def soak_case(out_path, rank, world, comm, total, every):
    """The low-latency exchange for `total` iterations; every `every` iterations the replicas of R
    are hashed against each other (a torn (value, tag) pair would leave one replica different)."""
    import hashlib
    import time
    inst = synth.random_instance(400, 300, 16, 0.08, total, 1e-5, seed=11)
    L, R = api.random_fill_LR(inst.users, inst.items, inst.features)
    plan = api.Plan(inst.users, inst.items, inst.features, inst.entries, inst.alpha, comm=comm)
    plan.set_factors(L, R)
    bad, done, t0 = 0, 0, time.time()
    while done < total:
        plan.iterate(every)
        done += every
        plan.get_factors(L, R)
        h = hashlib.sha256(R.tobytes()).hexdigest()
        hs = [None] * world
        dist.all_gather_object(hs, h)
        if len(set(hs)) != 1:
            bad += 1
    plan.close()
    comm.close()
    if rank == 0:
        line = "soak: %d ranks, %d iterations, replicas of R hashed every %d: %d mismatches, finite=%s, %.1f s" % (
            world, done, every, bad, bool(np.isfinite(R).all()), time.time() - t0)
        print(line, flush=True)
        with open(out_path, "w") as fh:
            fh.write(line + "\n")
    dist.destroy_process_group()
    if bad:
        sys.exit(1)


if __name__ == "__main__":
    main()
