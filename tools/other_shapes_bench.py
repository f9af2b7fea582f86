"""Perf points for the reference's remaining published shapes (deliverables/report-mpi.pdf p.4;
inputs missing upstream: synthetic stand-ins of the same users x items x features and nonzeros per
user) -- one JSON line per shape: time per iteration (CUDA events, 3 x 32
iterations from the reference's initial factors, best of 3), nonzero-updates/s, the per-kernel
times, the arg-max time, and the fraction of the HBM roofline of the SURVEY's compulsory bytes.

    python tools/other_shapes_bench.py > profiles/r02_other_shapes.jsonl
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from recommender_system_b200 import api, synth  # noqa: E402

SHAPES = [
    # name, users, items, F, (lo, hi) nonzeros per user
    ("inst1e6-100-700-1-3 (users scaled 1/10)", 100000, 100, 700, (1, 3)),
    ("inst1000-80000-20-10-1000", 1000, 80000, 20, (10, 1000)),
    ("inst20000-10000-40-2-50", 20000, 10000, 40, (2, 50)),
    ("inst50000-5000-100-2-5", 50000, 5000, 100, (2, 5)),
]


def main():
    peak = 6558.7
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as fh:
            peak = float(json.load(fh)["hbm_gbs"])
    except OSError:
        pass
    for name, U, I, F, (lo, hi) in SHAPES:
        inst = synth.inst60000_like(users=U, items=I, features=F, iters=32, alpha=0.0005, lo=lo, hi=hi, seed=777)
        L, R = api.random_fill_LR(U, I, F)
        with api.Plan(U, I, F, inst.entries, inst.alpha, ngpus=1) as plan:
            plan.set_factors(L, R)
            plan.save_factors()
            plan.iterate(32)
            plan.sync()
            best = 1e9
            for _ in range(3):
                plan.restore_factors()
                plan.timer_start()
                plan.iterate(32)
                best = min(best, plan.timer_stop() / 32)
            k = plan.profile_kernels(10)
            plan.recommend()
            rec = plan.last_recommend_ms()
            b_it = plan.bytes_per_iter()
        print(json.dumps({"shape": name, "users": U, "items": I, "features": F, "nnz": int(inst.nnz),
                          "us_per_iteration": best * 1e3, "nnz_iters_per_s": inst.nnz / (best * 1e-3),
                          "kernel_ms": k, "argmax_ms": rec, "algorithmic_bytes_per_iteration": b_it,
                          "hbm_roofline_frac": b_it / (best * 1e-3) / 1e9 / peak}), flush=True)


if __name__ == "__main__":
    main()
