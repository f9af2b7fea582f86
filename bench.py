#!/usr/bin/env python
"""bench.py -- nonzero-updates/s of the factorization loop (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A "step" is one complete factorization of the workload: T iterations of the
hot path (matFact.c:36-53) starting from the reference's initial factors.
`value` = nnz * T * K / device time of the K steps, inputs resident in HBM.
`e2e`   = the same metric through the public host-buffer call sequence
(plan create with the COO entries, factors H2D from pinned memory, T
iterations, factors + recommendations D2H), wall clock per step.
Under torchrun (N > 1) every rank owns a contiguous user range, R is
replicated and the partial R gradient is all-reduced each iteration; the
problem size is fixed, so scaling is "strong".
`cpu_baseline` (N = 1) and `--impl reference`: the reference's own CPU programs on all host
cores, on a bounded sample of the same workload -- its OpenMP matrix_factorization in process
(oracle/_ref/libref_omp.so) and its MPI program (oracle/_ref/matFact_mpi over the single-node MPI
stand-in oracle/mpi_shim), the faster of the two being the figure and the other kept beside it.

Workloads (BASELINE.json configs):
  inst60000  synthetic stand-in for inst60000-2000-200-10-20 (input blob missing
             upstream): U=60000 I=2000 F=200 nnz~0.9M, T=100.  DEFAULT: the one
             named shape whose factors (2 x 96 MB of L) do not fit the 126 MB L2.
  ml1m       synthetic stand-in for instML1M (blob missing): U=6040 I=3952 F=100
             nnz=1000209, T=100
  ml100k     the reference's samples/instML100k.in as is (T=3000, F=30)
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "nonzero-updates/sec (nnz*iters/s)"
UNIT = "nnz*iters/s"


def load_workload(name, iters_override=None):
    from recommender_system_b200 import synth
    from recommender_system_b200.io import read_instance
    if name == "inst60000":
        inst = synth.inst60000_like()
        desc = "inst60000-2000-200-10-20 shape, synthetic (seed 12346): U=60000 I=2000 F=200"
    elif name == "ml1m":
        inst = synth.ml_like()
        desc = "instML1M shape, synthetic (seed 12345): U=6040 I=3952 F=100"
    elif name == "ml100k":
        inst = read_instance(os.path.join(ROOT, "tests", "golden", "instML100k.in.gz"))
        desc = "samples/instML100k.in (reference sample): U=943 I=1682 F=30"
    else:
        raise SystemExit("unknown workload %r" % name)
    if iters_override:
        inst.iters = iters_override
    return inst, desc


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi style clock / throttle-reason samples during the timed region."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons = index, False, [], set()
        self.sm_max = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        self.stop_flag = True
        if self.is_alive():
            self.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def cpu_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def time_reference(inst, kind, budget_s, threads):
    """Loop-only nonzero-updates/s of the reference's own matrix_factorization
    (oracle/_ref/libref_<kind>.so, see oracle/ref_driver.c) on a bounded number of
    iterations of the same workload; the call's trailing mat2d_prod is removed by
    subtracting a 0-iteration call.  Falls back to the oracle port if the
    reference could not be compiled here."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    os.environ["OMP_NUM_THREADS"] = str(threads)
    os.environ.setdefault("OMP_PROC_BIND", "close")
    if kind == "omp" and pyoracle.ref_available(kind):
        # libgomp reads OMP_NUM_THREADS once; later calls in this process go through its API
        try:
            pyoracle._ref_lib(kind)
            ctypes.CDLL("libgomp.so.1").omp_set_num_threads(int(threads))
        except OSError:
            pass
    U, I, F = inst.users, inst.items, inst.features
    L0, R0 = pyoracle.init_LR(U, I, F)
    if pyoracle.ref_available(kind):
        def call(n):
            L, R = L0.copy(), R0.copy()
            return pyoracle.ref_factorize(kind, U, I, F, inst.entries, n, inst.alpha, L, R)
        label = "reference"
    else:
        def call(n):
            L, R = L0.copy(), R0.copy()
            t0 = time.perf_counter()
            pyoracle.factorize(U, I, F, inst.entries, n, inst.alpha, L, R)
            return time.perf_counter() - t0
        label, threads = "port", 1
    t_zero = min(call(0), call(0))
    t_two = call(2)
    per_iter = max((t_two - t_zero) / 2, 1e-6)
    n = int(max(2, min(inst.iters, budget_s / per_iter)))
    t_n = call(n)
    loop_s = max(t_n - t_zero, 1e-9)
    return {"value": inst.nnz * n / loop_s, "unit": UNIT, "cores": threads, "kind": label,
            "sample": "%d of %d iterations of the same workload, %s matrix_factorization "
                      "(%.2f s; 0-iteration call of %.2f s subtracted)"
                      % (n, inst.iters, "matFact-omp.c" if kind == "omp" else "matFact.c", loop_s, t_zero)}


_MPI_BODY = {}


def time_reference_mpi(inst, prog, nranks, threads, budget_s, n_iters=None):
    """_time_reference_mpi, or None (with a note on stderr) if the run fails: a baseline that cannot
    be taken must not take the bench line with it."""
    try:
        return _time_reference_mpi(inst, prog, nranks, threads, budget_s, n_iters)
    except Exception as exc:  # noqa: BLE001
        print("bench.py: %s baseline with %d ranks not taken: %s" % (prog, nranks, exc), file=sys.stderr)
        return None


def _time_reference_mpi(inst, prog, nranks, threads, budget_s, n_iters=None):
    """Loop-only nonzero-updates/s of the reference's distributed program (prog 'mpi' =
    matFact-mpi.c, 'hybrid' = matFact-hybrid.c, compiled unmodified against the single-node MPI
    stand-in oracle/mpi_shim -- no MPI on this box) with `nranks` processes x `threads` OpenMP
    threads, on a bounded number of iterations of the same workload written to a scratch input
    file; the time is the program's own MPI_Wtime bracket around matrix_factorization
    (matFact-mpi.c:647-653).  n_iters fixes the sample instead of sizing it from a 2-iteration run.
    None when oracle/_ref does not hold the program."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    import tempfile
    from recommender_system_b200.io import Instance, write_instance
    if not pyoracle.mpi_available(prog):
        return None
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "sample.in")

        def run(n):
            # the file is its iteration count on the first line, then a body that does not depend on it
            key = id(inst.entries)
            if key not in _MPI_BODY:
                write_instance(path, Instance(0, inst.alpha, inst.features, inst.users, inst.items, inst.entries))
                with open(path, "rb") as fh:
                    _MPI_BODY.clear()
                    _MPI_BODY[key] = fh.read().split(b"\n", 1)[1]
            with open(path, "wb") as fh:
                fh.write(b"%d\n" % n + _MPI_BODY[key])
            return pyoracle.run_ref_mpi(path, nranks, prog, threads, timeout=300)[1]
        if n_iters is None:
            per_iter = max(run(2) / 2, 1e-6)
            n = int(max(2, min(inst.iters, budget_s / per_iter)))
        else:
            n = int(n_iters)
        loop_s = max(run(n), 1e-9)
        if n_iters is None and loop_s < 0.4 * budget_s and n < inst.iters:
            # the 2-iteration estimate carries the start-up of the parallel regions: once more
            n = int(max(n, min(inst.iters, n * budget_s / loop_s)))
            loop_s = max(run(n), 1e-9)
    return {"value": inst.nnz * n / loop_s, "unit": UNIT, "cores": nranks * threads, "kind": "reference",
            "ranks": nranks, "threads_per_rank": threads, "iters_run": n,
            "sample": "%d of %d iterations of the same workload, %s over oracle/mpi_shim (single-node "
                      "shared-memory MPI stand-in), %.2f s by its own MPI_Wtime bracket"
                      % (n, inst.iters, "matFact-mpi.c" if prog == "mpi" else "matFact-hybrid.c", loop_s)}


def time_best_reference(inst, budget_s, threads):
    """The reference's CPU implementation with all the host threads: its OpenMP program and its MPI
    program (one rank per core, over oracle/mpi_shim) are both timed on a bounded sample; the
    faster one is the baseline, the other is kept beside it under its name."""
    omp = time_reference(inst, "omp", budget_s, threads)
    omp["program"] = "matFact-omp.c" if omp["kind"] == "reference" else "oracle port"
    mpi = time_reference_mpi(inst, "mpi", min(threads, 64), 1, budget_s) if omp["kind"] == "reference" else None
    if mpi is None:
        return omp
    mpi["program"] = "matFact-mpi.c over oracle/mpi_shim"
    best, other, name = (mpi, omp, "omp") if mpi["value"] > omp["value"] else (omp, mpi, "mpi")
    best[name] = other
    return best


def kernel_bytes(F, U_t, I_t, nnz):
    """Algorithmic (compulsory) HBM bytes of one ITERATION, SURVEY.md 8(d): every touched
    factor element read once and written once, A read once: 16*F*(U_t+I_t) + 12*nnz.  (The
    per-launch figures of the individual kernels come from the library,
    mfgpu_plan_kernel_bytes)"""
    it = 16 * F * (U_t + I_t) + 12 * nnz
    return None, None, it


def ncu_traffic(workload, kernel):
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None
    with open(p) as fh:
        return json.load(fh).get(workload, {}).get(kernel)


def make_config(desc, inst, n_gpus):
    """The `config` of the JSON line: the same dictionary in both arms (the driver compares them)."""
    return {"workload": desc, "iters": inst.iters, "nnz": int(inst.nnz), "alpha": inst.alpha,
            "sharding": "users" if n_gpus > 1 else "none",
            "l2": "factor buffers %.0f MB vs 126 MB L2; no flush (the loop re-reads its own output by "
                  "construction)" % (32e-6 * inst.features * (inst.users + inst.items))}


def run_reference_arm(args, rank):
    if rank != 0:
        return
    inst, desc = load_workload(args.workload, args.iters)
    threads = cpu_threads()
    budget = max(2.0, min(20.0, 150.0 / max(1, args.steps + args.warmup)))
    # warm-up doubles as the choice of program: the reference's OpenMP program on all host threads
    # against its MPI program with one rank per core (over oracle/mpi_shim); the faster one is timed
    probe = time_best_reference(inst, 2.0, threads)
    use_mpi = probe.get("program", "").startswith("matFact-mpi")
    n_mpi = int(max(2, min(inst.iters, budget * probe["value"] / max(1, inst.nnz)))) if use_mpi else None
    vals, last = [], None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        mpi_run = time_reference_mpi(inst, "mpi", probe["ranks"], 1, budget, n_iters=n_mpi) if use_mpi else None
        if mpi_run:
            last = mpi_run
            last["program"] = probe["program"]
        else:
            last = time_reference(inst, "omp", budget, threads)
            last["program"] = probe.get("program", "matFact-omp.c")
        vals.append(last["value"])
        if time.perf_counter() - t0 > 240:
            break
    other = "omp" if use_mpi else "mpi"
    if other in probe:
        last[other] = probe[other]
    v = float(np.mean(vals))
    last["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": len(vals), "warmup": args.warmup,
            "ms_per_step": 1e3 * inst.nnz * inst.iters / v, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic"
            if args.workload != "ml100k" else "reference sample file",
            "config": make_config(desc, inst, args.gpus),
            "note": "reference CPU implementation, all host threads: the faster of its OpenMP and MPI "
                    "programs (cpu_baseline.program); ms_per_step extrapolated to the full iteration count",
            "cpu_baseline": last,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


EXCHANGE = {0: "none", 1: "nccl allreduce of the R gradient + apply",
            2: "fused reduce+apply kernel over NVLink peer memory (CUDA IPC), rank order",
            3: "event-ordered reduce+apply kernel over NVLink peer memory, shard order (one process)",
            4: "chain kernels store gradient rows into the owning rank over NVLink peer memory; owner reduces in "
               "rank order, applies, stores R_new rows into every replica (fences + flag words)",
            5: "accumulation kernels store gradient elements as 16-byte (value, tag) pairs into the owning rank "
               "over NVLink peer memory; one kernel reduces the own items in rank order as pairs arrive, applies, "
               "stores R_new pairs into every replica and unpacks the others' (no fences, no flags)"}

# measured on this pool's B200 by tools/ubench_l2_gather.cu (profiles/r02_ubench_l2_gather.txt): rows of
# 1536 B gathered at random from a 3.2 MB table, 64 warps per SM -- what the L2 -> SM path delivers when
# enough loads are in flight (the loop kernels are bound by how many they can keep in flight)
L2_GATHER_PEAK_GBS = 27500.0


def expected_hash(workload, n_gpus, iters):
    """sha256 of the oracle's final L||R for this workload, iteration count and shard count, or None."""
    try:
        if workload == "ml100k":
            with open(os.path.join(ROOT, "tests", "golden", "expected.json")) as fh:
                rec = json.load(fh)["instML100k"]
            return rec["lr_sha256"] if n_gpus == 1 and iters == rec["iters"] else None
        with open(os.path.join(ROOT, "tests", "golden", "expected_shapes.json")) as fh:
            rec = json.load(fh)[workload]
        return rec["shards"].get(str(n_gpus)) if iters == rec["iters"] else None
    except (OSError, KeyError, ValueError):
        return None


class Ctx:
    """What every workload of one bench run shares: the process group, the communicator, the options."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from recommender_system_b200 import api
        self.torch, self.dist, self.api, self.args = torch, dist, api, args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if api.device_count() < 1:
            raise SystemExit("bench.py: no B200 visible and there is no CPU path")
        torch.cuda.set_device(self.local)
        self.comm = None
        if self.world > 1:
            if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
                os.environ["NCCL_DEBUG"] = "WARN"       # the version banner goes to stdout: one JSON line only
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            uid = torch.zeros(api.UID_BYTES, dtype=torch.uint8, device="cuda")
            if self.rank == 0:
                uid.copy_(torch.frombuffer(bytearray(api.Comm.unique_id()), dtype=torch.uint8))
            dist.broadcast(uid, 0)
            self.comm = api.Comm(self.world, self.rank, bytes(uid.cpu().numpy().tobytes()), self.local)
            self.cpu_group = dist.new_group(backend="gloo")     # waits that must not occupy a GPU
        self.n_gpus = self.world if self.world > 1 else args.gpus

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def host_barrier(self):
        """A barrier that waits on the host: no kernel spins on any GPU meanwhile."""
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier(group=self.cpu_group)

    def max_over_ranks(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def gather(self, obj):
        """obj of every rank, on every rank (pickled through the process group)."""
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out

    def close(self):
        if self.comm:
            self.comm.close()
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


def bench_workload(cx, name, steps, warmup, e2e_steps, headline):
    """One workload: the device-resident leg (`value`), the parity check of its factors against the
    oracle's hash on record, the per-kernel times, and the end-to-end leg.  Returns a dict (all
    ranks; rank 0's is printed)."""
    import hashlib
    torch, api, args = cx.torch, cx.api, cx.args
    inst, desc = load_workload(name, args.iters if headline else None)
    U, I, F, T = inst.users, inst.items, inst.features, inst.iters
    n_gpus, comm, rank = cx.n_gpus, cx.comm, cx.rank

    # the reference's host-side initialisation, kept in pinned memory for the e2e legs
    L0, R0 = api.random_fill_LR(U, I, F)
    Lp = torch.empty((U, F), dtype=torch.float64, pin_memory=True).numpy()
    Rp = torch.empty((I, F), dtype=torch.float64, pin_memory=True).numpy()
    Lo = torch.empty((U, F), dtype=torch.float64, pin_memory=True).numpy()
    Ro = torch.empty((I, F), dtype=torch.float64, pin_memory=True).numpy()
    Lp[...] = L0
    Rp[...] = R0
    del L0, R0

    def make_plan():
        pl = api.Plan(U, I, F, inst.entries, inst.alpha, ngpus=(1 if comm else args.gpus), comm=comm)
        for kv in args.option:
            k, v = kv.split("=")
            pl.set_option(k, int(v))
        return pl

    # ---- device-resident leg: `value` -------------------------------------------------
    plan = make_plan()
    xkind = plan.exchange_kind()
    plan.set_factors(Lp, Rp)
    plan.save_factors()
    for _ in range(warmup):
        plan.restore_factors()
        plan.iterate(T)
    plan.sync()
    cx.barrier()
    sampler = ClockSampler(cx.local)
    sampler.start()
    launches0 = plan.launch_count()
    plan.timer_start()
    for _ in range(steps):
        plan.restore_factors()
        plan.iterate(T)
    region_ms = plan.timer_stop()
    cx.barrier()
    clocks = sampler.result()
    region_ms = cx.max_over_ranks(region_ms)
    launches = plan.launch_count() - launches0
    value = inst.nnz * T * steps / (region_ms * 1e-3)

    # ---- parity of what was just timed: the factors of the last step against the oracle's hash --
    lo, hi = plan.user_range(0)
    single_process_multi = (not comm) and args.gpus > 1
    if single_process_multi:
        lo, hi = 0, U
    plan.get_factors(Lo, Ro)
    parts = cx.gather((lo, hi, Lo[lo:hi].copy() if comm else None, hashlib.sha256(Ro.tobytes()).hexdigest()))
    parity = None
    if rank == 0:
        if comm:
            for (a, b, rows, _) in parts:
                Lo[a:b] = rows
        h = hashlib.sha256()
        h.update(Lo.tobytes())
        h.update(Ro.tobytes())
        want = expected_hash(name, n_gpus, T)      # plan options never change a bit of the result
        parity = {"lr_sha256": h.hexdigest(), "expected": want,
                  "replicas_of_R_identical": len({p[3] for p in parts}) == 1,
                  "source": "oracle (tools/make_golden_shapes.py -> tests/golden/expected_shapes.json; "
                            "ml100k: tests/golden/expected.json), same cut, same order of the sum"}
        parity["checked"] = want is not None
        parity["ok"] = (h.hexdigest() == want) if want is not None else None
    del parts

    # per-kernel device times for the roofline (all ranks: the exchange is collective)
    k_ms = plan.profile_kernels(20)
    k_bytes = plan.kernel_bytes()
    g_bytes = plan.kernel_gather_bytes()
    lo0, hi0 = plan.user_range(0)
    e = inst.entries
    mine = e[(e["row"] >= lo0) & (e["row"] < hi0)]      # shard 0 (= this rank): the kernels timed above
    U_t, I_t = len(np.unique(mine["row"])), len(np.unique(mine["col"]))
    _, _, b_it = kernel_bytes(F, U_t, I_t, len(mine))
    peak, peak_src = peaks()
    loop_kernels = ["k_update_L", "k_dot", "k_chain_bulk", "k_chain_deep"]
    dom_name = max(loop_kernels, key=lambda k: k_ms[k])
    dom_bytes = k_bytes[dom_name]
    ach = dom_bytes / (k_ms[dom_name] * 1e-3) / 1e9 if k_ms[dom_name] > 0 else 0.0
    it_ms = region_ms / (steps * T)
    l2_ach = g_bytes[dom_name] / (k_ms[dom_name] * 1e-3) / 1e9 if k_ms[dom_name] > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": dom_name, "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, "traffic": ncu_traffic(name, dom_name),
                "peak_source": peak_src, "algorithmic_bytes_per_launch": dom_bytes,
                "kernel_ms": k_ms, "kernel_algorithmic_bytes": k_bytes,
                "kernel_names": "k_update_L = by-user pass (fused k_update_L[_res], or k_dot_units in split "
                                "mode); k_dot = dots of the long rows; k_chain_bulk = ordered accumulations "
                                "(k_chain, or k_acc in split mode); k_chain_deep = k_chain_ws",
                "what_binds": "the bytes a kernel keeps in flight from L2 (gathered factor rows, ~2000 cycles "
                              "under load), not HBM: see l2_gather and DESIGN.md section 4",
                "l2_gather": {"kernel": dom_name, "bytes_per_launch": g_bytes[dom_name], "achieved": l2_ach,
                              "peak": L2_GATHER_PEAK_GBS, "unit": "GB/s", "frac": l2_ach / L2_GATHER_PEAK_GBS,
                              "peak_source": "measured: tools/ubench_l2_gather.cu on this pool's B200, rows of 1536 B "
                                             "gathered from a 3.2 MB table (profiles/r02_ubench_l2_gather.txt)",
                              "kernel_gather_bytes": g_bytes},
                "iteration": {"algorithmic_bytes": b_it, "ms": it_ms,
                              "achieved": b_it / (it_ms * 1e-3) / 1e9,
                              "frac": b_it / (it_ms * 1e-3) / 1e9 / peak}}
    if n_gpus > 1:
        # kernel_ms.exchange_apply is the exchange launched on its own (for the low-latency
        # exchange together with a re-send of the contributions); what an iteration pays for the
        # exchange is what is left of it beside the loop kernels (the longer of each concurrent pair)
        beside = max(k_ms["k_update_L"], k_ms["k_dot"]) + max(k_ms["k_chain_bulk"], k_ms["k_chain_deep"])
        roofline["iteration"]["exchange_in_situ_ms"] = max(0.0, it_ms - beside)
    best = plan.recommend()
    rec_ms = plan.last_recommend_ms()
    plan.close()

    # ---- end-to-end leg: host buffers in, host buffers out, every step ------------------
    best2 = np.full(U, -2, dtype=np.int32)

    def e2e_step():
        t = [time.perf_counter()]
        pl = make_plan(); t.append(time.perf_counter())
        pl.set_factors(Lp, Rp); t.append(time.perf_counter())
        pl.iterate(T); t.append(time.perf_counter())
        pl.get_results(Lo, Ro, best2); t.append(time.perf_counter())
        pl.close(); t.append(time.perf_counter())
        if os.environ.get("MFGPU_TRACE"):       # host-side view; iterate returns before the device is done
            names = ["create", "set_factors", "iterate (enqueue)", "get_results: factors D2H beside the arg-max (+wait)", "close"]
            print("[bench rank %d] " % rank + ", ".join("%s %.1f ms" % (n, (b - a) * 1e3)
                                                        for n, a, b in zip(names, t, t[1:])), file=sys.stderr)

    e2e_step()
    cx.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    cx.barrier()
    e2e_s = cx.max_over_ranks(time.perf_counter() - t0)
    e2e_value = inst.nnz * T * e2e_steps / e2e_s
    nu = U if single_process_multi else hi0 - lo0
    nloc = inst.nnz if single_process_multi else len(mine)
    h2d = nloc * (4 + 8 + 4 + 4) + nloc * 4 + 4 * (2 * U + 2 + U_t + I_t) * 2 + 8 * F * (nu + I)
    d2h = 8 * F * (nu + I) + 4 * nu
    if not np.array_equal(best[lo0:hi0] if comm else best, best2[lo0:hi0] if comm else best2):
        raise SystemExit("bench.py: device-resident and end-to-end legs disagree")
    return {"inst": inst, "desc": desc, "value": value, "region_ms": region_ms, "steps": steps, "launches": launches,
            "clocks": clocks, "roofline": roofline, "parity": parity, "xkind": xkind,
            "argmax_ms": rec_ms,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "ms_per_step": 1e3 * e2e_s / e2e_steps}}


def bench_argmax_1e6(cx, ngpus=1):
    """BASELINE.json configs[4]: samples/inst1000-1e6-1000-1-3.in as it is -- its loop is 2e4
    nonzero-updates, its run time is the masked arg-max over L.R^T (2e12 separately rounded flops).
    R is 8 GB per buffer and is generated on the device(s).  On N GPUs it runs as ONE process
    driving the N GPUs (rank 0 of the bench; what `MATFACT_NGPUS=N matFact` does): the arg-max is
    sharded by ITEMS, every GPU scanning its block for all 1000 users.  Reports the arg-max time."""
    import hashlib
    from recommender_system_b200.io import format_output, read_instance
    api = cx.api
    path = os.path.join(ROOT, "tests", "golden", "inst1000-1e6-1000-1-3.in")
    if not os.path.exists(path):
        path += ".gz"
    inst = read_instance(path)
    t0 = time.perf_counter()
    with api.Plan(inst.users, inst.items, inst.features, inst.entries, inst.alpha, ngpus=ngpus) as plan:
        t1 = time.perf_counter()
        plan.init_factors()
        t2 = time.perf_counter()
        plan.iterate(inst.iters)
        loop_ms = plan.last_loop_ms()
        best = plan.recommend()
        ms = plan.last_recommend_ms()
        best = plan.recommend()
        ms = min(ms, plan.last_recommend_ms())
    flops = 2.0 * inst.users * inst.items * inst.features
    out = format_output(best)
    return {"workload": "samples/inst1000-1e6-1000-1-3.in: U=1000 I=1000000 F=1000 nnz=%d" % inst.nnz,
            "n_gpus": ngpus, "sharding": "items (one process, N GPUs)" if ngpus > 1 else "none",
            "argmax_ms": ms, "tflops_separately_rounded": flops / (ms * 1e-3) / 1e12,
            "frac_of_fp64_issue_roof": flops / (ms * 1e-3) / 18.6e12 / ngpus,
            "fp64_issue_roof": "64 FP64 lanes/clk/SM x 148 SMs x 1.965 GHz = 18.6e12 multiplies or adds per second "
                               "(no FMA: the reference rounds the product and the sum separately)",
            "loop_ms": loop_ms, "plan_create_s": t1 - t0, "init_factors_s": t2 - t1,
            "stdout_md5": hashlib.md5(out.encode()).hexdigest(),
            "stdout_md5_serial_reference": "cbb0f9797c19e495f70281dfbba2862d"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="inst60000", choices=["inst60000", "ml1m", "ml100k"])
    ap.add_argument("--iters", type=int, default=None, help="override the workload's iteration count")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="only the headline workload")
    ap.add_argument("--option", action="append", default=[], help="plan option name=value")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    cx = Ctx(args)
    r = bench_workload(cx, args.workload, args.steps, args.warmup, max(3, min(args.steps, 10)), True)
    inst = r["inst"]

    # the other named configurations, measured in the same run (BASELINE.json configs[1..4])
    extra = {}
    if not args.no_extra and not args.option:
        for name in ("ml1m", "ml100k", "inst60000"):
            if name == args.workload:
                continue
            x = bench_workload(cx, name, 3, 3, 3, False)
            xi = x["inst"]
            extra[name] = {"workload": x["desc"], "iters": xi.iters, "nnz": int(xi.nnz),
                           "value": x["value"], "unit": UNIT, "ms_per_step": x["region_ms"] / x["steps"],
                           "us_per_iteration": 1e3 * x["region_ms"] / (x["steps"] * xi.iters),
                           "e2e": x["e2e"], "argmax_ms": x["argmax_ms"], "clocks": x["clocks"],
                           "parity": x["parity"],
                           "roofline": {k: x["roofline"][k] for k in ("kernel", "achieved", "frac", "kernel_ms", "iteration", "l2_gather")}}
        # rank 0 alone drives all N GPUs for this one; the other ranks wait on the host, GPUs idle
        cx.host_barrier()
        if rank == 0:
            try:
                extra["inst1e6_argmax"] = bench_argmax_1e6(cx, cx.n_gpus)
            except Exception as exc:  # noqa: BLE001  (e.g. not enough host memory on the box)
                extra["inst1e6_argmax"] = {"skipped": str(exc)}
        cx.host_barrier()

    cpu = None
    if rank == 0 and cx.n_gpus == 1 and not args.no_cpu_baseline:
        cpu = time_best_reference(inst, 10.0, cpu_threads())
        if args.workload == "ml100k":
            # the reference's serial and OpenMP programs over the thread counts of its own report
            # (deliverables/report-omp.pdf Table 2: 1/2/4/8), on this box's host cores; cheap on this
            # input (its trailing mat2d_prod is 0.1 GFLOP), ~2.5 s each
            sweep = {"serial": time_reference(inst, "serial", 2.5, 1)["value"]}
            t = 1
            while t <= cpu_threads():
                sweep["omp_%d" % t] = time_reference(inst, "omp", 2.5, t)["value"]
                t *= 2
            t = 1
            while t <= min(cpu_threads(), 64):
                rr = time_reference_mpi(inst, "mpi", t, 1, 2.5)
                if rr:
                    sweep["mpi_%d" % t] = rr["value"]
                t *= 2
            if cpu_threads() >= 4:
                half = 1
                while half * 2 * 2 <= min(cpu_threads(), 64):
                    half *= 2
                rr = time_reference_mpi(inst, "hybrid", half, 2, 2.5)
                if rr:
                    sweep["hybrid_%dx2" % half] = rr["value"]
            cpu["by_threads"] = sweep
    n_gpus = cx.n_gpus
    cx.close()
    if rank != 0:
        return
    parity = r["parity"] or {}
    line = {"metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["region_ms"] / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic" if args.workload != "ml100k" else "reference sample file",
            "config": make_config(r["desc"], inst, n_gpus),
            "exchange": EXCHANGE[r["xkind"]],
            "parity_checked": bool(parity.get("checked") and parity.get("ok")), "parity": parity,
            "roofline": r["roofline"], "cpu_baseline": cpu, "e2e": r["e2e"],
            "gpu_launches": int(r["launches"]), "clocks": r["clocks"], "argmax_ms": r["argmax_ms"],
            "workloads": extra}
    if parity.get("checked") and not parity.get("ok"):
        print(json.dumps(line), flush=True)
        raise SystemExit("bench.py: the factors of the timed leg differ from the oracle's hash on record")
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
