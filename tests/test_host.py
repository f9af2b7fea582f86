"""CPU-side checks: the C-ABI library loads and exports every symbol the header
declares, the host logic (pattern builder, shard split, mask of print_output,
initialisation sequence, file format, CLI errors) behaves like the reference,
and the product path fails loudly without a GPU.  No compute call needs a GPU."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, gold_path, load_instance
from recommender_system_b200 import io as mfio
from recommender_system_b200 import synth
from recommender_system_b200.io import ENTRY_DTYPE, Instance


def test_library_exports_every_declared_symbol(built_lib):
    api = built_lib
    hdr = open(os.path.join(ROOT, "include", "matfact_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(mfgpu_[a-z_A-Z0-9]+)\s*\(", hdr)))
    assert len(declared) >= 25
    lib = api.lib()
    for name in declared:
        assert hasattr(lib, name), name
    assert sorted(api.EXPORTS) == declared
    nm = subprocess.run(["nm", "-D", "--defined-only", api.LIB_PATH], capture_output=True, text=True).stdout
    for name in declared:
        assert re.search(r"\bT %s\b" % name, nm), name
    assert b"sm_100a" in api.lib().mfgpu_version()


def test_no_cpu_fallback(built_lib):
    """Without a B200 every compute entry must fail with MFGPU_ENODEV (-3)."""
    api = built_lib
    if api.device_count() > 0:
        pytest.skip("a GPU is visible")
    inst = load_instance("inst0")
    L, R = api.random_fill_LR(inst.users, inst.items, inst.features)
    with pytest.raises(api.MatFactError) as ei:
        api.matrix_factorization(L, R, inst.entries, 1, inst.alpha)
    assert ei.value.code == -3 and "no CPU path" in str(ei.value)
    with pytest.raises(api.MatFactError):
        api.recommend(L, R, inst.entries)
    with pytest.raises(api.MatFactError):
        api.Plan(inst.users, inst.items, inst.features, inst.entries, inst.alpha)
    p = subprocess.run([os.path.join(ROOT, "bin", "matFact"), gold_path("inst0.in")],
                       capture_output=True, text=True)
    assert p.returncode == 255 and p.stdout == "" and p.stderr.startswith("Error: ")


def test_product_does_not_touch_the_oracle():
    """oracle/ is test infrastructure: nothing under the package, the Makefile's
    product targets or the header may name it."""
    pkg = os.path.join(ROOT, "recommender_system_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".c", ".h", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert "pyoracle" not in src and "liboracle" not in src and "oracle/" not in src, f
    assert "oracle" not in open(os.path.join(ROOT, "include", "matfact_b200.h")).read()


def test_cli_errors_match_reference(built_lib, tmp_path):
    """matFact.c:63-67 and util.c:7-34: messages on stderr, exit(-1) -> status 255."""
    exe = os.path.join(ROOT, "bin", "matFact")
    p = subprocess.run([exe], capture_output=True, text=True)
    assert p.returncode == 255
    assert p.stderr == "Run ./matFact.out fileError: Missing input file name.\n"
    p = subprocess.run([exe, str(tmp_path / "nope.in")], capture_output=True, text=True)
    assert p.returncode == 255 and p.stderr == "Error: Unable to open input file.\n"
    cases = {"x": "Error in int argument.", "3 y": "Error in double argument.",
             "3 0.1 z": "Error in int argument.", "3 0.1 2 4 5": "Error in multiple int argument.",
             "3 0.1 2 4 5 2 0 0 1.0 1 q 2.0": "Error in non-zero entry."}
    for text, msg in cases.items():
        f = tmp_path / "bad.in"
        f.write_text(text)
        p = subprocess.run([exe, str(f)], capture_output=True, text=True)
        assert p.returncode == 255 and p.stderr == "Error: %s\n" % msg and p.stdout == "", text


def test_random_fill_matches_oracle_and_known_answers(built_lib, oracle):
    api = built_lib
    for (u, i, f) in [(3, 5, 2), (30, 40, 10), (7, 1, 33)]:
        L, R = api.random_fill_LR(u, i, f)
        L0, R0 = oracle.init_LR(u, i, f)
        assert np.array_equal(L, L0) and np.array_equal(R, R0)
    L, _ = api.random_fill_LR(3, 5, 2)
    assert abs(L[0, 0] - 0.420094) < 1e-6      # samples/inst0.mats:6
    assert L[0, 0] == 1804289383 / 2147483647.0 / 2


def test_split_users_balances_nonzeros(built_lib):
    api = built_lib
    inst = load_instance("instML100k")
    for n in (1, 2, 4, 8):
        cut = api.split_users(inst.users, inst.entries, n)
        assert cut[0] == 0 and cut[-1] == inst.users and np.all(np.diff(cut) >= 0)
        rows = inst.entries["row"]
        per = [int(((rows >= cut[s]) & (rows < cut[s + 1])).sum()) for s in range(n)]
        assert sum(per) == inst.nnz
        assert max(per) <= inst.nnz / n + 737 + 1   # off by at most the longest row
    # no nonzeros: the reference's equal-count BLOCK_LOW split (mpiutil.h:8)
    cut = api.split_users(10, np.zeros(0, dtype=ENTRY_DTYPE), 4)
    assert list(cut) == [0, 2, 5, 7, 10]


def _cursor_mask(users, items, entries):
    """print_output's cursor, literally (matFact.c:14-22)."""
    aix, cnt, first = 0, np.zeros(users, int), np.full(users, -1)
    n = len(entries)
    for i in range(users):
        for j in range(items):
            if aix < n and entries[aix]["row"] == i and entries[aix]["col"] == j:
                aix += 1
                cnt[i] += 1
            elif first[i] < 0:
                first[i] = j
    return cnt, first


def test_mask_follows_the_reference_cursor(built_lib):
    api = built_lib
    rng = np.random.default_rng(1)
    inst = synth.random_instance(12, 17, 2, 0.4, 1, 0.1, seed=3)
    e = inst.entries
    variants = [e, e[::-1].copy(), np.concatenate([e[:20], e[10:30]]), e[rng.permutation(len(e))],
                np.concatenate([e[:15], e[40:50], e[15:40]]), e[:0]]
    full = synth.random_instance(3, 4, 2, 1.0, 1, 0.1).entries
    variants.append(full)
    for v in variants:
        users, items = (3, 4) if v is full else (12, 17)
        cnt, first = api.mask_stats(users, items, v)
        rc, rf = _cursor_mask(users, items, v)
        assert np.array_equal(cnt, rc) and np.array_equal(first, rf)


def test_bad_entries_are_rejected(built_lib):
    api = built_lib
    e = np.zeros(1, dtype=ENTRY_DTYPE)
    e["row"] = 5
    with pytest.raises(api.MatFactError) as ei:
        api.split_users(3, e, 2)
    assert ei.value.code == -1


def test_file_format_round_trip(tmp_path):
    inst = synth.random_instance(9, 11, 3, 0.3, 7, 0.25, seed=8)
    for name in ("a.in", "a.in.gz"):
        p = tmp_path / name
        mfio.write_instance(p, inst)
        back = mfio.read_instance(p)
        assert (back.iters, back.alpha, back.features, back.users, back.items) == (7, 0.25, 3, 9, 11)
        assert np.array_equal(back.entries, inst.entries)
    # token based like fscanf: line breaks carry no meaning (util.c:12-34)
    one_line = mfio.parse_instance(" ".join(open(tmp_path / "a.in").read().split()))
    assert np.array_equal(one_line.entries, inst.entries)
    assert mfio.format_output(np.array([3, -1, 0])) == "3\n0\n"


def test_synthetic_shapes():
    a = synth.inst60000_like(users=3000, items=2000)
    rc = np.bincount(a.entries["row"], minlength=a.users)
    assert rc.min() >= 10 and rc.max() <= 20 and a.features == 200
    key = a.entries["row"].astype(np.int64) * a.items + a.entries["col"]
    assert np.all(np.diff(key) > 0)                    # (row, col)-sorted, no duplicates
    b = synth.ml_like(users=600, items=400, nnz=30000, max_row=300)
    assert b.nnz == 30000
    rc = np.bincount(b.entries["row"], minlength=b.users)
    assert rc.min() >= 20 and rc.max() <= 300
    key = b.entries["row"].astype(np.int64) * b.items + b.entries["col"]
    assert np.all(np.diff(key) > 0)
    assert set(np.unique(b.entries["value"])) <= {1.0, 2.0, 3.0, 4.0, 5.0}


def test_threaded_pattern_build_is_deterministic(tmp_path):
    """host/mf_prep.c splits its two passes over threads and can build into a caller's workspace:
    every array must equal the one-thread build, also for unsorted / duplicated entry lists
    (tests/prep_check.c)."""
    exe = str(tmp_path / "prep_check")
    subprocess.run(["gcc", "-O2", "-Wall", "-Wextra", "-I" + os.path.join(ROOT, "include"),
                    "-I" + os.path.join(ROOT, "recommender_system_b200", "host"), "-o", exe,
                    os.path.join(ROOT, "tests", "prep_check.c"),
                    os.path.join(ROOT, "recommender_system_b200", "host", "mf_prep.c"), "-lpthread"], check=True)
    p = subprocess.run([exe], capture_output=True, text=True)
    assert p.returncode == 0 and p.stdout.strip().endswith("OK"), p.stdout


def test_host_ingest_matches_the_c_library(tmp_path):
    """host/mf_ingest.c: the lock-free restatement of glibc's random() equals
    srandom()/random() draw for draw, mf_random_fill_LR equals the reference's init sequence bit for
    bit, and the buffer parser classifies 200 000 fuzzed inputs exactly like the reference's fscanf
    sequence and returns the same values (tests/ingest_check.c)."""
    exe = str(tmp_path / "ingest_check")
    subprocess.run(["gcc", "-O2", "-Wall", "-Wextra", "-I" + os.path.join(ROOT, "include"),
                    "-I" + os.path.join(ROOT, "recommender_system_b200", "host"), "-o", exe,
                    os.path.join(ROOT, "tests", "ingest_check.c"),
                    os.path.join(ROOT, "recommender_system_b200", "host", "mf_ingest.c"), "-lpthread"], check=True)
    p = subprocess.run([exe], capture_output=True, text=True)
    assert p.returncode == 0 and p.stdout.strip().endswith("OK"), p.stdout[-2000:]


def test_library_parser_reads_the_samples_and_reports_reference_errors(built_lib, tmp_path):
    """mfgpu_read_instance through the C ABI: equal to the Python reader on the reference's
    samples; the reference's die() messages for broken files (util.c:12-34, matFact.c:69-70)."""
    import gzip
    api = built_lib
    for name in ("inst30-40-10-2-10.in", "inst0.in", "inst1.in", "inst2.in"):
        path = os.path.join(ROOT, "tests", "golden", name)
        inst = mfio.read_instance(path)
        it, al, f, u, m, ent = api.read_instance(path)
        assert (it, al, f, u, m) == (inst.iters, inst.alpha, inst.features, inst.users, inst.items)
        assert np.array_equal(ent, inst.entries)
    big = tmp_path / "ml100k.in"
    with gzip.open(os.path.join(ROOT, "tests", "golden", "instML100k.in.gz"), "rb") as fh:
        big.write_bytes(fh.read())
    inst = mfio.read_instance(str(big))
    it, al, f, u, m, ent = api.read_instance(str(big))
    assert (it, al, f, u, m) == (inst.iters, inst.alpha, inst.features, inst.users, inst.items)
    assert np.array_equal(ent, inst.entries)
    cases = {"": "Error in int argument.", "10 x": "Error in double argument.",
             "10 0.1 y": "Error in int argument.", "10 0.1 2 3 4": "Error in multiple int argument.",
             "10 0.1 2 3 4 2 0 0 1.0 1": "Error in non-zero entry."}
    for text, msg in cases.items():
        bad = tmp_path / "bad.in"
        bad.write_text(text)
        with pytest.raises(api.MatFactError, match=msg.replace(".", r"\.")):
            api.read_instance(str(bad))
    with pytest.raises(api.MatFactError, match="Unable to open input file"):
        api.read_instance(str(tmp_path / "missing.in"))
    assert api.selftest_random(0, 200000) == -1 and api.selftest_random(7, 1000) == -1


def test_reference_main_with_the_patch_has_no_cpu_path():
    """oracle/_ref/matFact_patched = the reference's own main with the two calls at the seam
    replaced by mfgpu_run (oracle/integration_patch.sed).  Without a B200 it must stop the way the
    reference stops on an error (util.c:7-10: "Error: ...", exit status 255), not compute on the CPU."""
    import subprocess
    exe = os.path.join(ROOT, "oracle", "_ref", "matFact_patched")
    if not os.path.exists(exe):
        pytest.skip("not built (no reference checkout at build time)")
    from recommender_system_b200 import api
    if api.device_count() > 0:
        pytest.skip("a B200 is visible: covered by the gpu test of the same program")
    p = subprocess.run([exe, os.path.join(ROOT, "tests", "golden", "inst0.in")], capture_output=True, text=True)
    assert p.returncode == 255 and p.stderr.startswith("Error: ") and p.stdout == ""
