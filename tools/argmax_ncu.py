"""One masked arg-max on the device-generated initial factors of a workload (for ncu):
    python tools/argmax_ncu.py inst1e6|inst60000"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recommender_system_b200 import api, synth  # noqa: E402
from recommender_system_b200.io import read_instance  # noqa: E402

w = sys.argv[1] if len(sys.argv) > 1 else "inst1e6"
inst = read_instance(os.path.join(ROOT, "tests", "golden", "inst1000-1e6-1000-1-3.in")) if w == "inst1e6" else synth.inst60000_like()
with api.Plan(inst.users, inst.items, inst.features, inst.entries, inst.alpha, ngpus=1) as plan:
    plan.init_factors()
    plan.recommend()
    print(w, "argmax_ms", plan.last_recommend_ms())
