"""Kernel-time sweeps on one GPU (development aid, not part of the bench contract).

    python tools/tune.py <users> <items> <features> [name=value,name=value ...]
    python tools/tune.py ml:<users>:<nnz> <items> <features> [...]      (MovieLens-like rows)

Builds an inst60000-like instance of the given shape (10..20 nonzeros per user, seed 12346),
and for every option set given (comma-separated plan options; '-' = defaults) prints the
per-kernel times of mfgpu_plan_profile_kernels and the time per iteration of 3 x 64 iterations.
users = 7500 is what one rank of eight holds of the inst60000 shape.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from recommender_system_b200 import api, synth  # noqa: E402


def main():
    I, F = int(sys.argv[2]), int(sys.argv[3])
    sets = sys.argv[4:] or ["-"]
    if sys.argv[1].startswith("ml:"):
        _, u, n = sys.argv[1].split(":")
        U = int(u)
        inst = synth.ml_like(users=U, items=I, nnz=int(n), features=F)
    else:
        U = int(sys.argv[1])
        inst = synth.inst60000_like(users=U, items=I, features=F, lo=int(os.environ.get('TUNE_LO', 10)),
                                    hi=int(os.environ.get('TUNE_HI', 20)))
    L, R = api.random_fill_LR(U, I, F)
    print("U=%d I=%d F=%d nnz=%d" % (U, I, F, inst.nnz), flush=True)
    for opts in sets:
        with api.Plan(U, I, F, inst.entries, inst.alpha, ngpus=1) as plan:
            if opts != "-":
                for kv in opts.split(","):
                    k, v = kv.split("=")
                    plan.set_option(k, int(v))
            plan.set_factors(L, R)
            plan.save_factors()
            plan.iterate(64)
            plan.sync()
            best = 1e9
            for _ in range(3):
                plan.restore_factors()
                plan.timer_start()
                plan.iterate(64)
                best = min(best, plan.timer_stop() / 64)
            k = plan.profile_kernels(20)
            print("%-40s iter %7.1f us | %s" % (opts, best * 1e3, "  ".join(
                "%s %6.1f" % (n, 1e3 * t) for n, t in k.items() if t > 1e-3)), flush=True)


if __name__ == "__main__":
    main()
