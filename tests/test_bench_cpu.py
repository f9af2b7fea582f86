"""bench.py's reference arm (`--impl reference`) runs on host cores only, so its contract can be
checked here: one JSON line from rank 0 with the keys the driver reads, nothing from the other
ranks, and the bench's choice between the reference's OpenMP and MPI programs recorded in
cpu_baseline.  (The GPU arm needs a B200: covered by the driver's own bench run.)"""
import json
import os
import subprocess
import sys

from conftest import ROOT


def _run(env_extra, *args):
    env = dict(os.environ, **env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "ml100k",
                           "--iters", "30", "--steps", "1", "--warmup", "1", *args],
                          capture_output=True, text=True, env=env, timeout=600)


def test_reference_arm_line(oracle):
    oracle.build(ref=True)      # keeps a prebuilt oracle/_ref when the reference checkout is absent
    p = _run({})
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "nnz*iters/s" and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["gpu_launches"] == 0
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    c = d["cpu_baseline"]
    assert c["value"] == d["value"] and c["cores"] >= 1 and c["kind"] in ("reference", "port")
    assert "iterations of the same workload" in c["sample"]
    if c["kind"] == "reference" and oracle.mpi_available("mpi"):
        # both of the reference's parallel programs were timed; the faster one is the line
        other = "omp" if c["program"].startswith("matFact-mpi") else "mpi"
        assert other in c and c[other]["value"] > 0


def test_reference_arm_other_ranks_stay_silent():
    p = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, "--gpus", "2")
    assert p.returncode == 0 and p.stdout.strip() == ""
