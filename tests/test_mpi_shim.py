"""oracle/mpi_shim -- the single-node MPI stand-in -- and, through it, the
reference's own distributed programs.

1. tests/mpi_shim_check.c: the stand-in's semantics on their own (long messages, chunked
   collectives, the order of the sum, derived type + user operation, Gatherv, grid,
   sub-communicators, two threads in collectives at once), at 1..8 ranks.
2. oracle/_ref/matFact_mpi and matFact_hybrid (matFact-mpi.c / matFact-hybrid.c compiled
   UNMODIFIED against the stand-in by oracle/Makefile) reproduce the reference's golden
   stdout at several rank counts -- which pins the stand-in against the programs it exists
   for.  Skipped when oracle/_ref was not built (no reference checkout on this machine).

CPU only; nothing here touches the product library.
"""
import gzip
import os
import shutil
import subprocess

import pytest

from conftest import GOLD, ROOT, gold_path, load_out

SHIM = os.path.join(ROOT, "oracle", "mpi_shim")


@pytest.fixture(scope="module")
def launcher(oracle):
    return oracle.mpirun_shim()


@pytest.fixture(scope="module")
def shim_check(tmp_path_factory, launcher):
    exe = str(tmp_path_factory.mktemp("shim") / "mpi_shim_check")
    subprocess.run(["gcc", "-O2", "-Wall", "-Wextra", "-Werror", "-I", SHIM, "-o", exe,
                    os.path.join(ROOT, "tests", "mpi_shim_check.c"), os.path.join(SHIM, "mpi_shim.c"),
                    "-lpthread"], check=True)
    return exe


@pytest.mark.parametrize("nranks", [1, 2, 3, 4, 6, 8])
def test_shim_semantics(launcher, shim_check, nranks):
    p = subprocess.run([launcher, "-np", str(nranks), shim_check], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    assert p.stdout.strip() == "OK %d" % nranks


def test_launcher_reports_a_failing_rank(tmp_path, launcher):
    """One rank exits like the reference's die() (util.c:7-10: status 255) while the others wait
    in a barrier: the job ends with that status instead of hanging."""
    src = tmp_path / "die.c"
    src.write_text('#include <mpi.h>\n#include <stdlib.h>\nint main(int c, char **v) { int r; MPI_Init(&c, &v);'
                   ' MPI_Comm_rank(MPI_COMM_WORLD, &r); if (r == 1) exit(-1); MPI_Barrier(MPI_COMM_WORLD);'
                   ' MPI_Finalize(); return 0; }\n')
    exe = str(tmp_path / "die")
    subprocess.run(["gcc", "-O2", "-I", SHIM, "-o", exe, str(src), os.path.join(SHIM, "mpi_shim.c")], check=True)
    p = subprocess.run([launcher, "-np", "4", exe], capture_output=True, timeout=60)
    assert p.returncode == 255
    # without the launcher a program is a job of one rank
    p = subprocess.run([exe], capture_output=True, timeout=60)
    assert p.returncode == 0


def test_ranks_die_with_a_killed_launcher(tmp_path, launcher):
    """A caller's timeout kills the launcher with SIGKILL (subprocess.run(timeout=...)): the ranks,
    spinning in a barrier that will never complete, must not outlive it."""
    import signal
    import time
    src = tmp_path / "hang.c"
    src.write_text('#include <mpi.h>\n#include <unistd.h>\nint main(int c, char **v) { int r; MPI_Init(&c, &v);'
                   ' MPI_Comm_rank(MPI_COMM_WORLD, &r); if (r == 0) sleep(600); MPI_Barrier(MPI_COMM_WORLD);'
                   ' MPI_Finalize(); return 0; }\n')
    exe = str(tmp_path / "hang_for_shim_test")
    subprocess.run(["gcc", "-O2", "-I", SHIM, "-o", exe, str(src), os.path.join(SHIM, "mpi_shim.c")], check=True)
    p = subprocess.Popen([launcher, "-np", "3", exe])
    time.sleep(1.0)
    p.send_signal(signal.SIGKILL)
    p.wait()
    deadline = time.time() + 10
    while time.time() < deadline:
        alive = [d for d in os.listdir("/proc") if d.isdigit() and _cmdline(d).startswith(exe)]
        if not alive:
            break
        time.sleep(0.2)
    assert not alive


def _cmdline(pid):
    try:
        with open("/proc/%s/cmdline" % pid, "rb") as fh:
            return fh.read().split(b"\0")[0].decode(errors="replace")
    except OSError:
        return ""


def _plain_input(sample, tmp_path):
    p = gold_path(sample + ".in")
    if not p.endswith(".gz"):
        return p
    out = tmp_path / (sample + ".in")
    with gzip.open(p, "rb") as src, open(out, "wb") as dst:
        shutil.copyfileobj(src, dst)
    return str(out)


CASES = [  # (sample, program, ranks, threads per rank)
    ("inst0", "mpi", 1, 1), ("inst0", "mpi", 2, 1), ("inst0", "mpi", 8, 1),   # more ranks than rows
    ("inst1", "mpi", 3, 1), ("inst2", "mpi", 6, 1),
    ("inst30-40-10-2-10", "mpi", 4, 1), ("inst30-40-10-2-10", "mpi", 7, 1),
    ("inst500-500-20-2-100", "mpi", 8, 1),
    ("inst200-10000-50-100-300", "mpi", 4, 1),    # items >> users: the grid is transposed (mpiutil.c:60-83)
    ("inst600-10000-10-40-400", "mpi", 6, 1),
    ("inst30-40-10-2-10", "hybrid", 2, 2), ("inst500-500-20-2-100", "hybrid", 4, 2),
    ("inst200-10000-50-100-300", "hybrid", 2, 3),
]


@pytest.mark.parametrize("sample,prog,nranks,threads", CASES)
def test_reference_distributed_programs_reproduce_the_goldens(oracle, tmp_path, sample, prog, nranks, threads):
    if not oracle.mpi_available(prog):
        pytest.skip("oracle/_ref/matFact_%s not built on this machine" % prog)
    out, took = oracle.run_ref_mpi(_plain_input(sample, tmp_path), nranks, prog, threads, timeout=600)
    assert out == load_out(sample)
    assert took is not None and took >= 0.0
