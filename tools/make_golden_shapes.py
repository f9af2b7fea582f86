#!/usr/bin/env python
"""Record the oracle's final L||R for the two headline shapes whose upstream inputs are missing
(synthetic stand-ins, recommender_system_b200/synth.py), at the benchmark's full iteration count
and for the user cuts the GPU path uses at 1, 2, 4 and 8 shards:

    python tools/make_golden_shapes.py         # ~1-2 minutes of CPU (oracle/, test infrastructure)

Writes tests/golden/expected_shapes.json:
    {workload: {"iters": T, "nnz": .., "shards": {"1": sha256, "2": .., "4": .., "8": ..},
                "cut": {"2": [...], ...}, "best_sha256": {"1": ..}}}
sha256 is over the raw doubles of L (users x F) followed by R (items x F).  shards "1" is the
serial order of matFact.c:41-53; N > 1 is oracle_factorize_sharded with the cut of
mfgpu_split_users (R_new = R_old + ((G_0 + G_1) + ...), the order every peer-memory exchange of
the library keeps).  bench.py compares the factors of its device-resident leg with these
("parity_checked"), tests/test_gpu_parity.py the full-size runs.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pyoracle  # noqa: E402
from recommender_system_b200 import api, synth  # noqa: E402


def sha(L, R):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(L).tobytes())
    h.update(np.ascontiguousarray(R).tobytes())
    return h.hexdigest()


def main():
    pyoracle.build(ref=False)
    out = {}
    for name, inst in (("inst60000", synth.inst60000_like()), ("ml1m", synth.ml_like())):
        U, I, F, T = inst.users, inst.items, inst.features, inst.iters
        rec = {"iters": T, "nnz": int(inst.nnz), "users": U, "items": I, "features": F, "alpha": inst.alpha,
               "shards": {}, "cut": {}, "best_sha256": {}}
        for n in (1, 2, 4, 8):
            t0 = time.time()
            L, R = pyoracle.init_LR(U, I, F)
            if n == 1:
                pyoracle.factorize(U, I, F, inst.entries, T, inst.alpha, L, R)
            else:
                cut = api.split_users(U, inst.entries, n)      # host-side helper of the library, no GPU
                rec["cut"][str(n)] = [int(c) for c in cut]
                pyoracle.factorize_sharded(U, I, F, inst.entries, T, inst.alpha, L, R, cut)
            assert np.isfinite(L).all() and np.isfinite(R).all(), "the synthetic header diverges"
            rec["shards"][str(n)] = sha(L, R)
            if n == 1:
                best = pyoracle.recommend(U, I, F, inst.entries, L, R)
                rec["best_sha256"]["1"] = hashlib.sha256(np.ascontiguousarray(best, dtype=np.int32).tobytes()).hexdigest()
            print(name, n, rec["shards"][str(n)][:16], "%.1f s" % (time.time() - t0), flush=True)
        out[name] = rec
    path = os.path.join(ROOT, "tests", "golden", "expected_shapes.json")
    with open(path, "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)
    print("wrote", path)


if __name__ == "__main__":
    main()
