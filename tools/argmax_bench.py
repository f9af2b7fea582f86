"""Masked arg-max phase (mat2d_prod + print_output, matFact.c:55,10-27) timed on one GPU.

    python tools/argmax_bench.py [inst1e6|inst60000|ml1m] ...

For each workload: the reference's initial factors, its iteration count of the loop, then the
recommendation; prints one JSON line with the device time of the arg-max launches, the
2*U*I*F flops they stand for and the rate (FP64, multiply and add rounded separately, so the
pipe's ceiling is half its FMA figure).  inst1e6 = samples/inst1000-1e6-1000-1-3.in (kept under
tests/golden), the configuration SURVEY.md 8(d) asks to be reported this way."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recommender_system_b200 import api, synth  # noqa: E402
from recommender_system_b200.io import read_instance  # noqa: E402


def main():
    for w in sys.argv[1:] or ["inst1e6"]:
        if w == "inst1e6":
            inst = read_instance(os.path.join(ROOT, "tests", "golden", "inst1000-1e6-1000-1-3.in"))
        elif w == "ml1m":
            inst = synth.ml_like()
        else:
            inst = synth.inst60000_like()
        U, I, F = inst.users, inst.items, inst.features
        t0 = time.perf_counter()
        L, R = api.random_fill_LR(U, I, F)
        t_init = time.perf_counter() - t0
        with api.Plan(U, I, F, inst.entries, inst.alpha, ngpus=1) as plan:
            t0 = time.perf_counter()
            plan.set_factors(L, R)
            plan.sync()
            t_h2d = time.perf_counter() - t0
            plan.iterate(inst.iters)
            loop_ms = plan.last_loop_ms()
            plan.recommend()
            ms = min(plan.last_recommend_ms() for _ in range(3) if plan.recommend() is not None)
        flops = 2.0 * U * I * F
        print(json.dumps({"workload": w, "users": U, "items": I, "features": F, "nnz": inst.nnz,
                          "init_s": round(t_init, 3), "h2d_s": round(t_h2d, 3), "loop_ms": round(loop_ms, 3),
                          "argmax_ms": round(ms, 3), "flops": flops,
                          "tflops_fp64_no_fma": round(flops / (ms * 1e-3) / 1e12, 2)}), flush=True)


if __name__ == "__main__":
    main()
