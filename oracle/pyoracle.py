"""ctypes binding of the CPU oracle -- TEST INFRASTRUCTURE ONLY.

May be imported from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never from the product package.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle.so")
CLI = os.path.join(HERE, "_build", "matfact_oracle")
REF_DIR = os.path.join(HERE, "_ref")

_lib = None
# the CPUs this process may use, before any OpenMP runtime binds the calling thread
try:
    _ALL_CPUS = os.sched_getaffinity(0)
except AttributeError:
    _ALL_CPUS = None


def build(ref: bool = True) -> None:
    """Compile the restatement (and oracle/_ref when /root/reference exists)."""
    targets = ["all"] + (["ref"] if ref else [])
    subprocess.run(["make", "-C", HERE] + targets, check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build(ref=False)
        L = ctypes.CDLL(LIB)
        dp = ctypes.POINTER(ctypes.c_double)
        ip = ctypes.POINTER(ctypes.c_int)
        vp = ctypes.c_void_p
        i = ctypes.c_int
        L.oracle_init_LR.argtypes = [i, i, i, dp, dp]
        L.oracle_init_LR.restype = None
        L.oracle_factorize.argtypes = [i, i, i, i, vp, i, ctypes.c_double, dp, dp]
        L.oracle_factorize.restype = i
        L.oracle_factorize_sharded.argtypes = [i, i, i, i, vp, i, ctypes.c_double, dp, dp, i, ip]
        L.oracle_factorize_sharded.restype = i
        L.oracle_recommend.argtypes = [i, i, i, i, vp, dp, dp, ip]
        L.oracle_recommend.restype = i
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def init_LR(users, items, features):
    L = np.empty((users, features), dtype=np.float64)
    R = np.empty((items, features), dtype=np.float64)
    lib().oracle_init_LR(users, items, features, _dp(L), _dp(R))
    return L, R


def factorize(users, items, features, entries, iters, alpha, L, R):
    entries = np.ascontiguousarray(entries)
    assert entries.dtype.itemsize == 16 and L.flags.c_contiguous and R.flags.c_contiguous
    rc = lib().oracle_factorize(users, items, features, len(entries), entries.ctypes.data,
                                iters, alpha, _dp(L), _dp(R))
    if rc:
        raise MemoryError("oracle_factorize")


def factorize_sharded(users, items, features, entries, iters, alpha, L, R, shard_lo):
    entries = np.ascontiguousarray(entries)
    lo = np.ascontiguousarray(shard_lo, dtype=np.int32)
    rc = lib().oracle_factorize_sharded(users, items, features, len(entries), entries.ctypes.data,
                                        iters, alpha, _dp(L), _dp(R), len(lo) - 1,
                                        lo.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
    if rc:
        raise MemoryError("oracle_factorize_sharded")


def recommend(users, items, features, entries, L, R):
    entries = np.ascontiguousarray(entries)
    best = np.empty(users, dtype=np.int32)
    rc = lib().oracle_recommend(users, items, features, len(entries), entries.ctypes.data,
                                _dp(L), _dp(R), best.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
    if rc:
        raise MemoryError("oracle_recommend")
    return best


def run_instance(inst):
    """Whole reference program on an Instance: returns (L, R, best)."""
    L, R = init_LR(inst.users, inst.items, inst.features)
    factorize(inst.users, inst.items, inst.features, inst.entries, inst.iters, inst.alpha, L, R)
    best = recommend(inst.users, inst.items, inst.features, inst.entries, L, R)
    return L, R, best


# ---- the UNMODIFIED reference, in process (oracle/_ref/libref_*.so) -------------------

_ref_libs = {}


def ref_available(kind="serial"):
    return os.path.exists(os.path.join(REF_DIR, "libref_%s.so" % kind))


def _ref_lib(kind):
    if kind not in _ref_libs:
        L = ctypes.CDLL(os.path.join(REF_DIR, "libref_%s.so" % kind))
        i = ctypes.c_int
        L.ref_factorize.argtypes = [i, i, i, i, ctypes.c_void_p, i, ctypes.c_double,
                                    ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
        L.ref_factorize.restype = ctypes.c_double
        _ref_libs[kind] = L
    return _ref_libs[kind]


def ref_factorize(kind, users, items, features, entries, iters, alpha, L, R):
    """The reference's own matrix_factorization (kind: 'serial' = matFact.c:29-59,
    'omp' = matFact-omp.c:35-144; OMP_NUM_THREADS picks its threads) on L, R in
    place.  Returns the seconds of the call (includes its trailing mat2d_prod)."""
    entries = np.ascontiguousarray(entries)
    s = _ref_lib(kind).ref_factorize(users, items, features, len(entries), entries.ctypes.data,
                                     iters, alpha, _dp(L), _dp(R))
    if s < 0:
        raise MemoryError("ref_factorize")
    return s


def mpi_available(prog="mpi"):
    """True when oracle/_ref holds the reference's distributed program (matFact-mpi.c or
    matFact-hybrid.c built against oracle/mpi_shim)."""
    return os.path.exists(os.path.join(REF_DIR, "matFact_%s" % prog))


def mpirun_shim():
    """Path of the launcher of the MPI stand-in (oracle/mpi_shim), compiled on demand."""
    p = os.path.join(HERE, "_build", "mpirun_shim")
    if not os.path.exists(p):
        build(ref=False)
    return p


def run_ref_mpi(path, nranks, prog="mpi", threads=1, timeout=None):
    """Run the UNMODIFIED reference program (prog: 'mpi' = matFact-mpi.c, 'hybrid' =
    matFact-hybrid.c) on the input file `path` with `nranks` processes under
    oracle/_build/mpirun_shim (threads = OMP_NUM_THREADS of each rank).  Returns
    (recommendation lines as the serial program would print them, seconds its own
    MPI_Wtime bracket reports for matrix_factorization -- matFact-mpi.c:647-653).
    The programs interleave diagnostics on stdout; the recommendations are the lines
    that are a bare integer."""
    # OMP_PROC_BIND must not be inherited: libgomp would bind the initial thread of EVERY rank to
    # the first place (all ranks on one core)
    env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="false")
    # ... and neither must the CPU mask of this thread, which an in-process OpenMP run with
    # OMP_PROC_BIND may have narrowed to one core (children inherit it)
    mask = os.sched_getaffinity(0) if _ALL_CPUS else None
    if _ALL_CPUS:
        os.sched_setaffinity(0, _ALL_CPUS)
    try:
        p = subprocess.run([mpirun_shim(), "-np", str(nranks),
                            os.path.join(REF_DIR, "matFact_%s" % prog), str(path)],
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env, timeout=timeout, text=True)
    finally:
        if _ALL_CPUS:
            os.sched_setaffinity(0, mask)
    if p.returncode != 0:
        raise RuntimeError("matFact_%s -np %d exited %d: %s" % (prog, nranks, p.returncode, p.stderr[-400:]))
    out, took = [], None
    for line in p.stdout.splitlines():
        if line.isdigit():
            out.append(line)
        elif line.startswith("Matrix factorization took "):
            took = float(line.split()[3])
    return "".join(s + "\n" for s in out), took
