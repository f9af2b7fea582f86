"""`.mats`-style trace of a run on the GPU path (debugging aid for parity).

    python tools/mats_trace.py <file.in> [--iters 0,1,2,3,4] [--no-b]

Prints what samples/inst0.mats holds -- "Initial matrix A/L/R/B", then "Iter=<n>" / "Matrix L/R/B"
after each listed iteration (0-based, as in the sample traces), then "Final:" -- with %f, R as
features x items and B = L.R^T summed over k in order (mat2d.c:100-113), from the factors the
library holds on the device after each step.  `diff` against a trace of the reference or of
oracle/_build/matfact_oracle shows the first iteration and element where two runs part.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from recommender_system_b200 import api  # noqa: E402
from recommender_system_b200.io import read_instance  # noqa: E402


def product(L, R):
    """B[i][j] = sum over k, from 0.0, in order; no FMA (numpy multiplies and adds separately)."""
    B = np.zeros((L.shape[0], R.shape[0]))
    for k in range(L.shape[1]):
        B = B + np.outer(L[:, k], R[:, k])
    return B


def fmt(name, M):
    return name + "\n" + "".join(" ".join("%f" % x for x in row) + " \n" for row in M)


def section(L, R, prefix, with_b):
    out = fmt(prefix + "L", L) + fmt(prefix + "R", R.T)
    if with_b:
        out += fmt(prefix + "B", product(L, R))
    return out


def trace(inst, iters, with_b=True):
    """The trace as one string; iters = the 0-based iterations after which a section is printed."""
    U, I, F = inst.users, inst.items, inst.features
    A = np.zeros((U, I))
    A[inst.entries["row"], inst.entries["col"]] = inst.entries["value"]
    L, R = api.random_fill_LR(U, I, F)
    out = [fmt("Initial matrix A", A), section(L, R, "Initial matrix ", with_b)]
    with api.Plan(U, I, F, inst.entries, inst.alpha) as plan:
        plan.set_factors(L, R)
        done = 0
        for it in sorted(i for i in iters if 0 <= i < inst.iters):
            plan.iterate(it + 1 - done)
            done = it + 1
            plan.get_factors(L, R)
            out.append("Iter=%d\n" % it + section(L, R, "Matrix ", with_b))
        if inst.iters > done:
            plan.iterate(inst.iters - done)
        plan.get_factors(L, R)
    out.append("Final:\n" + section(L, R, "Matrix ", with_b))
    return "".join(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("file")
    ap.add_argument("--iters", default="0,1,2,3,4")
    ap.add_argument("--no-b", action="store_true", help="leave B out (it is users x items)")
    args = ap.parse_args()
    iters = [int(x) for x in args.iters.split(",") if x != ""]
    sys.stdout.write(trace(read_instance(args.file), iters, not args.no_b))


if __name__ == "__main__":
    main()
