"""mfgpu_recommend_noalloc (include/matfact_b200.h): the masked arg-max on caller-supplied DEVICE
buffers with its scratch in a built-in device array -- no cudaMalloc inside the call.  No
counterpart in the reference; checked here against
the existing entry point mfgpu_recommend (host buffers, pool scratch) on the same inputs:
integer output, tolerance 0, on two calls in a row (the scratch array keeps the previous call's
contents).  Collected last (test_zz_*) so a failure here hides no other test under pytest -x."""
import numpy as np
import pytest

from recommender_system_b200 import synth

pytestmark = pytest.mark.gpu


def _device_inputs(torch, api, inst, L, R):
    users, items, F = inst.users, inst.items, inst.features
    ld = (F + 1) & ~1
    e = inst.entries
    cnt, ff = api.mask_stats(users, items, e)
    mask_ptr = np.zeros(users + 1, dtype=np.int32)
    np.cumsum(cnt, out=mask_ptr[1:])
    assert mask_ptr[-1] == len(e)                      # sorted, duplicate-free: every entry is matched
    Lp = np.zeros((users, ld)); Lp[:, :F] = L
    Rp = np.zeros((items, ld)); Rp[:, :F] = R
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return ld, t(Lp), t(Rp), t(mask_ptr), t(e["col"].astype(np.int32)), t(ff.astype(np.int32))


@pytest.mark.parametrize("shape", [(300, 1500, 37, 0.05), (70, 130, 8, 0.5)])
def test_noalloc_matches_recommend(gpu, shape):
    import torch
    users, items, F, dens = shape
    inst = synth.random_instance(users, items, F, dens, 0, 0.0, seed=11)
    rng = np.random.default_rng(5)
    dev = torch.device("cuda:0")
    d_best = torch.full((users,), -7, dtype=torch.int32, device=dev)
    with torch.cuda.device(0):
        for call in range(2):                               # two calls in a row, different factors
            L = rng.standard_normal((users, F))
            R = rng.standard_normal((items, F))
            if call == 1:
                R[3] = R[1]                                 # an exact tie: lowest index wins
                L[2, 0] = np.nan                            # a NaN row: the first unrated item stays
            ref = gpu.recommend(L, R, inst.entries)
            ld, dL, dR, dptr, dcol, dff = _device_inputs(torch, gpu, inst, L, R)
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated()
            gpu.recommend_noalloc(users, items, F, ld, dL.data_ptr(), dR.data_ptr(), dptr.data_ptr(),
                                  dcol.data_ptr(), dff.data_ptr(), d_best.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated() == before
            assert np.array_equal(d_best.cpu().numpy(), ref), "call %d differs from mfgpu_recommend" % call


def test_noalloc_rejects_oversize(gpu):
    import torch
    with torch.cuda.device(0):
        x = torch.zeros(16, dtype=torch.float64, device="cuda:0")
        with pytest.raises(gpu.MatFactError) as ei:
            gpu.recommend_noalloc(300000, 4, 2, 2, x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(),
                                  x.data_ptr(), x.data_ptr(), 0)
        assert ei.value.code == -1 and "scratch" in str(ei.value)
