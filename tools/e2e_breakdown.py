import time, sys, numpy as np
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
from recommender_system_b200 import api, synth
import torch
w = sys.argv[1] if len(sys.argv) > 1 else 'inst60000'
inst = synth.inst60000_like() if w == 'inst60000' else synth.ml_like()
L, R = api.random_fill_LR(inst.users, inst.items, inst.features)
Lp = torch.from_numpy(L).pin_memory().numpy(); Rp = torch.from_numpy(R).pin_memory().numpy()
Lo = torch.empty_like(torch.from_numpy(L)).pin_memory().numpy(); Ro = torch.empty_like(torch.from_numpy(R)).pin_memory().numpy()
best = np.zeros(inst.users, dtype=np.int32)
for rep in range(4):
    t = [time.perf_counter()]
    pl = api.Plan(inst.users, inst.items, inst.features, inst.entries, inst.alpha); t.append(time.perf_counter())
    pl.set_factors(Lp, Rp); pl.sync(); t.append(time.perf_counter())
    pl.iterate(inst.iters); pl.sync(); t.append(time.perf_counter())
    pl.get_factors(Lo, Ro); t.append(time.perf_counter())
    pl.recommend(best); t.append(time.perf_counter())
    pl.close(); t.append(time.perf_counter())
    names = ['create', 'set_factors', 'iterate', 'get_factors', 'recommend', 'close']
    print(w, ' '.join('%s %.1f' % (n, (b - a) * 1e3) for n, a, b in zip(names, t, t[1:])), 'total %.1f ms' % ((t[-1] - t[0]) * 1e3))
