"""ctypes binding of libmatfact_b200.so (include/matfact_b200.h).

Host-side mirror of the reference's two calls at the seam,
``matrix_factorization(B, L, R, entries, nnz, iters, alpha)`` (matFact.c:29-59)
and ``print_output(B, entries)`` (matFact.c:10-27), with the same argument
meaning: L is users x features, R is items x features (transposed factor),
both row-major float64 updated IN PLACE; entries use the reference's 16-byte
``non_zero_entry`` layout.  B is never materialised.

There is no CPU path here: if the shared library is missing or no B200 is
visible every call raises ``MatFactError``.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from .io import ENTRY_DTYPE, Instance, format_output

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MFGPU_LIB") or os.path.join(HERE, "libmatfact_b200.so")   # MFGPU_LIB: a variant build (tools/build_variant.sh)

EXPORTS = [
    "mfgpu_device_count", "mfgpu_last_error", "mfgpu_version",
    "mfgpu_factorize", "mfgpu_recommend", "mfgpu_run",
    "mfgpu_plan_create", "mfgpu_plan_create_rank", "mfgpu_plan_destroy",
    "mfgpu_plan_set_factors", "mfgpu_plan_init_factors", "mfgpu_plan_get_factors", "mfgpu_plan_get_results", "mfgpu_plan_iterate", "mfgpu_plan_sync",
    "mfgpu_plan_last_loop_ms", "mfgpu_plan_recommend", "mfgpu_plan_last_recommend_ms",
    "mfgpu_plan_user_range", "mfgpu_plan_launch_count", "mfgpu_plan_exchange_kind", "mfgpu_plan_bytes_per_iter",
    "mfgpu_plan_set_option", "mfgpu_plan_save_factors", "mfgpu_plan_restore_factors",
    "mfgpu_plan_profile_kernels", "mfgpu_plan_kernel_bytes", "mfgpu_plan_kernel_gather_bytes", "mfgpu_plan_timer_start", "mfgpu_plan_timer_stop",
    "mfgpu_comm_unique_id", "mfgpu_comm_create", "mfgpu_comm_destroy",
    "mfgpu_split_users", "mfgpu_random_fill_LR", "mfgpu_mask_stats",
    "mfgpu_read_instance", "mfgpu_read_error", "mfgpu_free", "mfgpu_selftest_random",
    "mfgpu_exchange_layout", "mfgpu_recommend_noalloc",
]

UID_BYTES = 128
KERNELS = ["k_update_L", "k_dot", "k_chain_bulk", "k_chain_deep", "exchange_apply"]


class MatFactError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("matfact_b200 error %d: %s" % (code, msg))
        self.code = code


_lib = None


def lib():
    """Load the C-ABI library (fails loudly when it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MatFactError(-3, "%s not built (run `make` or __graft_entry__.build())" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    i, d, vp, cp = ctypes.c_int, ctypes.c_double, ctypes.c_void_p, ctypes.c_char_p
    pp = ctypes.POINTER(ctypes.c_void_p)
    L.mfgpu_device_count.restype = i
    L.mfgpu_last_error.restype = cp
    L.mfgpu_version.restype = cp
    L.mfgpu_factorize.argtypes = [i, i, i, i, vp, i, d, vp, vp, i]
    L.mfgpu_recommend.argtypes = [i, i, i, i, vp, vp, vp, vp, i]
    L.mfgpu_run.argtypes = [i, i, i, i, vp, i, d, vp, vp, vp, i]
    L.mfgpu_plan_create.argtypes = [pp, i, i, i, i, vp, d, i]
    L.mfgpu_plan_create_rank.argtypes = [pp, i, i, i, i, vp, d, vp]
    L.mfgpu_plan_destroy.argtypes = [vp]
    L.mfgpu_plan_destroy.restype = None
    L.mfgpu_plan_set_factors.argtypes = [vp, vp, vp]
    L.mfgpu_plan_init_factors.argtypes = [vp]
    L.mfgpu_plan_get_factors.argtypes = [vp, vp, vp]
    L.mfgpu_plan_get_results.argtypes = [vp, vp, vp, vp]
    L.mfgpu_plan_iterate.argtypes = [vp, i]
    L.mfgpu_plan_sync.argtypes = [vp]
    L.mfgpu_plan_last_loop_ms.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    L.mfgpu_plan_recommend.argtypes = [vp, vp]
    L.mfgpu_plan_last_recommend_ms.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    L.mfgpu_plan_user_range.argtypes = [vp, i, ctypes.POINTER(i), ctypes.POINTER(i)]
    L.mfgpu_plan_launch_count.argtypes = [vp]
    L.mfgpu_plan_launch_count.restype = ctypes.c_longlong
    L.mfgpu_plan_exchange_kind.argtypes = [vp]
    L.mfgpu_plan_exchange_kind.restype = ctypes.c_int
    L.mfgpu_plan_bytes_per_iter.argtypes = [vp]
    L.mfgpu_plan_bytes_per_iter.restype = ctypes.c_longlong
    L.mfgpu_plan_set_option.argtypes = [vp, cp, i]
    L.mfgpu_plan_save_factors.argtypes = [vp]
    L.mfgpu_plan_restore_factors.argtypes = [vp]
    L.mfgpu_plan_profile_kernels.argtypes = [vp, i, ctypes.POINTER(ctypes.c_float)]
    L.mfgpu_plan_kernel_bytes.argtypes = [vp, ctypes.POINTER(ctypes.c_longlong)]
    L.mfgpu_plan_kernel_gather_bytes.argtypes = [vp, ctypes.POINTER(ctypes.c_longlong)]
    L.mfgpu_plan_timer_start.argtypes = [vp]
    L.mfgpu_plan_timer_stop.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    L.mfgpu_comm_unique_id.argtypes = [vp]
    L.mfgpu_comm_create.argtypes = [pp, i, i, vp, i]
    L.mfgpu_comm_destroy.argtypes = [vp]
    L.mfgpu_comm_destroy.restype = None
    L.mfgpu_split_users.argtypes = [i, i, vp, i, vp]
    L.mfgpu_random_fill_LR.argtypes = [i, i, i, vp, vp]
    L.mfgpu_random_fill_LR.restype = None
    L.mfgpu_read_instance.argtypes = [cp, ctypes.POINTER(i), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(i),
                                      ctypes.POINTER(i), ctypes.POINTER(i), ctypes.POINTER(i),
                                      ctypes.POINTER(vp)]
    L.mfgpu_read_error.argtypes = [i]
    L.mfgpu_read_error.restype = cp
    L.mfgpu_free.argtypes = [vp]
    L.mfgpu_free.restype = None
    L.mfgpu_selftest_random.argtypes = [ctypes.c_uint, ctypes.c_long]
    L.mfgpu_selftest_random.restype = ctypes.c_long
    L.mfgpu_exchange_layout.argtypes = [i, i, i, vp, i, vp, vp, ctypes.POINTER(i)]
    L.mfgpu_mask_stats.argtypes = [i, i, i, vp, vp, vp]
    L.mfgpu_recommend_noalloc.argtypes = [i, i, i, i, vp, vp, vp, vp, vp, vp, vp]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise MatFactError(rc, lib().mfgpu_last_error().decode())


def _entries(entries):
    e = np.ascontiguousarray(entries, dtype=ENTRY_DTYPE)
    return e


def _f64(a, shape):
    if not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous
            and a.shape == tuple(shape)):
        raise ValueError("expected a C-contiguous float64 array of shape %r" % (tuple(shape),))
    return a


def device_count() -> int:
    return lib().mfgpu_device_count()


def read_instance(path):
    """The library's parser (the reference's fscanf sequence, matFact.c:72-110): returns
    (iters, alpha, features, users, items, entries) or raises MatFactError carrying the
    reference's die() message for the first error."""
    it, f, u, m, nz = (ctypes.c_int() for _ in range(5))
    al = ctypes.c_double()
    ent = ctypes.c_void_p()
    rc = lib().mfgpu_read_instance(str(path).encode(), ctypes.byref(it), ctypes.byref(al), ctypes.byref(f),
                                   ctypes.byref(u), ctypes.byref(m), ctypes.byref(nz), ctypes.byref(ent))
    if rc != 0:
        raise MatFactError(rc, lib().mfgpu_read_error(rc).decode())
    try:
        buf = (ctypes.c_char * (16 * nz.value)).from_address(ent.value) if nz.value else b""
        entries = np.frombuffer(bytes(buf), dtype=ENTRY_DTYPE).copy()
    finally:
        lib().mfgpu_free(ent)
    return it.value, al.value, f.value, u.value, m.value, entries


def exchange_layout(users, items, entries, nranks):
    """Rank-mode exchange layout (host/mf_prep.c): (item_slot[items], contrib[n_touched],
    slots_per_rank).  Touched item of slot t is reduced by rank t // slots_per_rank; bit r of
    contrib[t] = rank r holds a nonzero of it."""
    e = _entries(entries)
    slot = np.empty(max(items, 1), dtype=np.int32)
    contrib = np.zeros(max(items, 1), dtype=np.int32)
    per = ctypes.c_int()
    n = lib().mfgpu_exchange_layout(users, items, len(e), e.ctypes.data, nranks, slot.ctypes.data,
                                    contrib.ctypes.data, ctypes.byref(per))
    if n < 0:
        raise MatFactError(n, lib().mfgpu_last_error().decode())
    return slot[:items], contrib[:n], per.value


def selftest_random(seed=0, n=1_000_000):
    """Index of the first draw where the library's restatement of glibc's random() differs from
    the C library's own after srandom(seed); -1 if the first n agree."""
    return lib().mfgpu_selftest_random(seed, n)


def random_fill_LR(users, items, features):
    """mat2d_random_fill_LR + mat2d_transpose (mat2d.c:61-72,115-124): host side."""
    L = np.empty((users, features), dtype=np.float64)
    R = np.empty((items, features), dtype=np.float64)
    lib().mfgpu_random_fill_LR(users, items, features, L.ctypes.data, R.ctypes.data)
    return L, R


def split_users(users, entries, nshards):
    e = _entries(entries)
    cut = np.zeros(nshards + 1, dtype=np.int32)
    _check(lib().mfgpu_split_users(users, len(e), e.ctypes.data, nshards, cut.ctypes.data))
    return cut


def mask_stats(users, items, entries):
    e = _entries(entries)
    cnt = np.zeros(users, dtype=np.int32)
    ff = np.zeros(users, dtype=np.int32)
    rc = lib().mfgpu_mask_stats(users, items, len(e), e.ctypes.data, cnt.ctypes.data, ff.ctypes.data)
    if rc < 0:
        _check(rc)
    return cnt, ff


def recommend_noalloc(users, items, features, ld, dL, dR, d_mask_ptr, d_mask_col, d_first_free, d_best,
                      stream=0):
    """mfgpu_recommend_noalloc: every pointer argument is a DEVICE address (an int); nothing is
    allocated by the call; the work is enqueued on `stream` (a cudaStream_t as an int)."""
    _check(lib().mfgpu_recommend_noalloc(users, items, features, ld, dL, dR, d_mask_ptr, d_mask_col,
                                         d_first_free, d_best, stream))


def matrix_factorization(L, R, entries, iters, alpha, ngpus=1):
    """matFact.c:29-53 on the GPU: updates L and R in place (host buffers)."""
    users, F = L.shape
    items = R.shape[0]
    _f64(L, (users, F)); _f64(R, (items, F))
    e = _entries(entries)
    _check(lib().mfgpu_factorize(users, items, F, len(e), e.ctypes.data, iters, alpha,
                                 L.ctypes.data, R.ctypes.data, ngpus))


def recommend(L, R, entries, ngpus=1):
    """mat2d_prod + print_output (matFact.c:55,10-27): item per user, -1 = no line."""
    users, F = L.shape
    items = R.shape[0]
    _f64(L, (users, F)); _f64(R, (items, F))
    e = _entries(entries)
    best = np.empty(users, dtype=np.int32)
    _check(lib().mfgpu_recommend(users, items, F, len(e), e.ctypes.data, L.ctypes.data,
                                 R.ctypes.data, best.ctypes.data, ngpus))
    return best


def print_output(L, R, entries, ngpus=1) -> str:
    """The text the reference's print_output would write to stdout."""
    return format_output(recommend(L, R, entries, ngpus))


def run(L, R, entries, iters, alpha, ngpus=1):
    """Both calls with the factors staying in HBM in between; returns best."""
    users, F = L.shape
    items = R.shape[0]
    _f64(L, (users, F)); _f64(R, (items, F))
    e = _entries(entries)
    best = np.empty(users, dtype=np.int32)
    _check(lib().mfgpu_run(users, items, F, len(e), e.ctypes.data, iters, alpha,
                           L.ctypes.data, R.ctypes.data, best.ctypes.data, ngpus))
    return best


def run_instance(inst: Instance, ngpus=1):
    """Whole program on an Instance: reference init, factorize, recommend."""
    L, R = random_fill_LR(inst.users, inst.items, inst.features)
    best = run(L, R, inst.entries, inst.iters, inst.alpha, ngpus)
    return L, R, best


class Comm:
    """One rank of a multi-process group (one process per GPU)."""

    def __init__(self, nranks, rank, uid: bytes, device):
        self._h = ctypes.c_void_p()
        buf = ctypes.create_string_buffer(bytes(uid), UID_BYTES)
        _check(lib().mfgpu_comm_create(ctypes.byref(self._h), nranks, rank, buf, device))
        self.nranks, self.rank, self.device = nranks, rank, device

    @staticmethod
    def unique_id() -> bytes:
        buf = ctypes.create_string_buffer(UID_BYTES)
        _check(lib().mfgpu_comm_unique_id(buf))
        return buf.raw

    def close(self):
        if self._h:
            lib().mfgpu_comm_destroy(self._h)
            self._h = ctypes.c_void_p()


class Plan:
    """The problem resident in HBM across calls (mfgpu_plan_*)."""

    def __init__(self, users, items, features, entries, alpha, ngpus=1, comm: Comm | None = None):
        self.users, self.items, self.features = users, items, features
        self._entries = _entries(entries)
        self.nnz = len(self._entries)
        self._h = ctypes.c_void_p()
        if comm is None:
            _check(lib().mfgpu_plan_create(ctypes.byref(self._h), users, items, features, self.nnz,
                                           self._entries.ctypes.data, alpha, ngpus))
        else:
            _check(lib().mfgpu_plan_create_rank(ctypes.byref(self._h), users, items, features, self.nnz,
                                                self._entries.ctypes.data, alpha, comm._h))

    def set_factors(self, L, R):
        _f64(L, (self.users, self.features)); _f64(R, (self.items, self.features))
        _check(lib().mfgpu_plan_set_factors(self._h, L.ctypes.data, R.ctypes.data))

    def init_factors(self):
        """The reference's initial L and R (srandom(0) sequence) generated on the device."""
        _check(lib().mfgpu_plan_init_factors(self._h))

    def get_factors(self, L, R):
        _f64(L, (self.users, self.features)); _f64(R, (self.items, self.features))
        _check(lib().mfgpu_plan_get_factors(self._h, L.ctypes.data, R.ctypes.data))

    def get_results(self, L, R, best=None):
        """get_factors + recommend in one call (the copies run beside the arg-max)."""
        if best is None:
            best = np.full(self.users, -2, dtype=np.int32)
        _f64(L, (self.users, self.features)); _f64(R, (self.items, self.features))
        _check(lib().mfgpu_plan_get_results(self._h, L.ctypes.data, R.ctypes.data, best.ctypes.data))
        return best

    def iterate(self, iters):
        _check(lib().mfgpu_plan_iterate(self._h, iters))

    def sync(self):
        _check(lib().mfgpu_plan_sync(self._h))

    def last_loop_ms(self) -> float:
        ms = ctypes.c_float()
        _check(lib().mfgpu_plan_last_loop_ms(self._h, ctypes.byref(ms)))
        return ms.value

    def recommend(self, best=None):
        if best is None:
            best = np.full(self.users, -2, dtype=np.int32)
        _check(lib().mfgpu_plan_recommend(self._h, best.ctypes.data))
        return best

    def last_recommend_ms(self) -> float:
        ms = ctypes.c_float()
        _check(lib().mfgpu_plan_last_recommend_ms(self._h, ctypes.byref(ms)))
        return ms.value

    def user_range(self, shard=0):
        lo, hi = ctypes.c_int(), ctypes.c_int()
        _check(lib().mfgpu_plan_user_range(self._h, shard, ctypes.byref(lo), ctypes.byref(hi)))
        return lo.value, hi.value

    def launch_count(self) -> int:
        return lib().mfgpu_plan_launch_count(self._h)

    def exchange_kind(self):
        """0 none, 1 NCCL all-reduce, 2 peer memory across processes (pull), 3 peer memory in one
        process, 4 peer memory across processes (push: chain kernels store into the reducing rank,
        fences + flags), 5 the same with (value, tag) pairs and no fences (low latency, default)."""
        return lib().mfgpu_plan_exchange_kind(self._h)

    def bytes_per_iter(self) -> int:
        return lib().mfgpu_plan_bytes_per_iter(self._h)

    def set_option(self, name, value):
        _check(lib().mfgpu_plan_set_option(self._h, name.encode(), int(value)))

    def timer_start(self):
        _check(lib().mfgpu_plan_timer_start(self._h))

    def timer_stop(self) -> float:
        ms = ctypes.c_float()
        _check(lib().mfgpu_plan_timer_stop(self._h, ctypes.byref(ms)))
        return ms.value

    def save_factors(self):
        _check(lib().mfgpu_plan_save_factors(self._h))

    def restore_factors(self):
        _check(lib().mfgpu_plan_restore_factors(self._h))

    def profile_kernels(self, reps=20):
        ms = (ctypes.c_float * 5)()
        _check(lib().mfgpu_plan_profile_kernels(self._h, reps, ms))
        return dict(zip(KERNELS, [float(x) for x in ms]))

    def kernel_bytes(self):
        b = (ctypes.c_longlong * 5)()
        _check(lib().mfgpu_plan_kernel_bytes(self._h, b))
        return dict(zip(KERNELS, [int(x) for x in b]))

    def kernel_gather_bytes(self):
        b = (ctypes.c_longlong * 5)()
        _check(lib().mfgpu_plan_kernel_gather_bytes(self._h, b))
        return dict(zip(KERNELS, [int(x) for x in b]))

    def close(self):
        if self._h:
            lib().mfgpu_plan_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
