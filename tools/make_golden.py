#!/usr/bin/env python
"""Regenerate tests/golden/ from the read-only reference checkout.

Run in the build container (needs /root/reference):

    python tools/make_golden.py            # all samples with a usable golden
    python tools/make_golden.py --quick    # skip the four >20 s samples

What it commits to tests/golden/:
  * samples/X.in, X.out, X.mats of the reference, verbatim (gzip for > 64 KiB);
    these are the reference's own golden vectors for this path;
  * expected.json: per sample the header, sha256 of the final L||R doubles as
    produced by the oracle, sha256 of the expected stdout, and a record that
    (a) oracle stdout == upstream .out and (b) the UNMODIFIED reference serial
    program built into oracle/_ref prints the same stdout.
The GPU box has no /root/reference; tests there read only these fixtures.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import gzip
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recommender_system_b200.io import read_golden_out, read_instance  # noqa: E402

REF = "/root/reference/samples"
GOLD = os.path.join(ROOT, "tests", "golden")
ORACLE = os.path.join(ROOT, "oracle", "_build", "matfact_oracle")
REFBIN = os.path.join(ROOT, "oracle", "_ref", "matFact_serial")

SAMPLES = [
    "inst0", "inst1", "inst2", "inst30-40-10-2-10", "inst500-500-20-2-100",
    "inst1000-1000-100-2-30", "inst200-10000-50-100-300", "inst400-50000-30-200-500",
    "inst600-10000-10-40-400", "inst50000-5000-100-2-5", "instML100k",
]
SLOW = {"inst400-50000-30-200-500", "inst600-10000-10-40-400", "inst50000-5000-100-2-5", "instML100k"}
NO_GOLDEN = ["inst1000-1e6-1000-1-3"]  # upstream .out is 0 bytes


def stash(name: str) -> str:
    src = os.path.join(REF, name)
    if os.path.getsize(src) > 64 * 1024:
        dst = os.path.join(GOLD, name + ".gz")
        with open(src, "rb") as f, gzip.GzipFile(dst, "wb", mtime=0) as g:
            shutil.copyfileobj(f, g)
    else:
        dst = os.path.join(GOLD, name)
        shutil.copyfile(src, dst)
        os.chmod(dst, 0o644)
    return os.path.basename(dst)


def run_one(sample: str, with_ref: bool, norec: bool = False):
    src = os.path.join(REF, sample + ".in")
    with tempfile.TemporaryDirectory() as td:
        dump = os.path.join(td, "lr.bin")
        cmd = [ORACLE, src, dump] + (["norec"] if norec else [])
        p = subprocess.run(cmd, capture_output=True, text=True, check=True)
        lr = hashlib.sha256(open(dump, "rb").read()).hexdigest()
        loop_s = float(p.stderr.split("loop_s")[1].split()[0])
    rec = {"lr_sha256": lr, "oracle_loop_s": round(loop_s, 3)}
    if not norec:
        rec["out_sha256"] = hashlib.sha256(p.stdout.encode()).hexdigest()
        gold = read_golden_out(os.path.join(REF, sample + ".out"))
        rec["oracle_matches_upstream_out"] = (p.stdout == gold)
        if with_ref:
            r = subprocess.run([REFBIN, src], capture_output=True, text=True, check=True)
            rec["ref_serial_matches_oracle"] = (r.stdout == p.stdout)
    return sample, rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    os.makedirs(GOLD, exist_ok=True)
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "all", "ref"], check=True,
                   stdout=subprocess.DEVNULL)

    path = os.path.join(GOLD, "expected.json")
    expected = json.load(open(path)) if os.path.exists(path) else {}
    todo = [s for s in SAMPLES if not (args.quick and s in SLOW)]
    with cf.ThreadPoolExecutor(max_workers=os.cpu_count()) as ex:
        futs = [ex.submit(run_one, s, True) for s in todo]
        futs += [ex.submit(run_one, s, False, True) for s in NO_GOLDEN]
        for f in cf.as_completed(futs):
            s, rec = f.result()
            inst = read_instance(os.path.join(REF, s + ".in"))
            rec.update(iters=inst.iters, alpha=inst.alpha, features=inst.features, users=inst.users,
                       items=inst.items, nnz=inst.nnz)
            rec["files"] = [stash(s + ".in")]
            if s not in NO_GOLDEN:
                rec["files"].append(stash(s + ".out"))
            if os.path.exists(os.path.join(REF, s + ".mats")):
                rec["files"].append(stash(s + ".mats"))
            old = expected.get(s, {})
            old.update(rec)
            expected[s] = old
            print(s, rec, flush=True)
    json.dump(expected, open(path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
