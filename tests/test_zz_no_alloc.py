"""GPU: the allocation-free entry point mpcb_value_grad_static (include/mpcb200.h) against the existing
mpcb_value_grad, through the C ABI, on two calls in a row.  Collected last on purpose (file name):
a failure here must not hide any other test.  Passes alone in a fresh process and after the
other tests: it uses its own handle and its own tensors, and the static scratch needs no initial contents.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-9  # the rollout / cost tolerance of tests/test_gpu_parity.py (the two entry points share one kernel)


def _inputs(mp, torch, B, N, seed, model=0):
    from mpcb200 import workloads as wl
    w = wl.tracking_batch(B, N, 100, seed=seed, model="pacejka" if model == 0 else "kin", random_U=True, d_min=0.3)
    dev = torch.device("cuda", torch.cuda.current_device())
    t = {k: torch.tensor(w[k], device=dev) for k in ("x0", "U0", "road", "road_idx")}
    return t


@pytest.mark.parametrize("model,N", [(0, 20), (1, 12)])
def test_value_grad_static_matches_value_grad_twice(model, N):
    import torch
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device (there is no CPU fallback to test instead)")
    import mpcb200 as mp
    B = 3000  # ragged: not a multiple of the CTA size
    prob = mp.BatchedMpc(N=N, S=100, model_id=model, road_mode=mp.ROAD_TABLE, n_roads=2)
    dev = prob.device
    sets = [_inputs(mp, torch, B, N, seed, model) for seed in (901, 902)]
    # reference results from the existing entry point (allocates its workspace on first use)
    refs = [prob.value_grad(t["x0"], t["U0"], t["road"], road_idx=t["road_idx"]) for t in sets]
    # outputs allocated up front: nothing below may allocate
    outs = [(torch.full((B,), float("nan"), dtype=torch.float64, device=dev),
             torch.full((B, 2 * N), float("nan"), dtype=torch.float64, device=dev)) for _ in sets]
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    alloc0, resv0 = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)
    l0 = prob.launch_count()
    # two calls in a row, different inputs: the second sees the first call's leftovers in the static scratch
    for t, (c, g) in zip(sets, outs):
        prob.value_grad_static(t["x0"], t["U0"], t["road"], c, g, road_idx=t["road_idx"])
    torch.cuda.synchronize()
    assert prob.launch_count() - l0 == 2          # one kernel per call
    assert torch.cuda.memory_allocated(dev) == alloc0 and torch.cuda.memory_reserved(dev) == resv0
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, f"device memory changed by {free0 - free1} bytes: something allocated"
    for (c_ref, g_ref), (c, g) in zip(refs, outs):
        c_ref, g_ref, c, g = (x.cpu().numpy() for x in (c_ref, g_ref, c, g))
        assert np.isfinite(c).all() and np.isfinite(g).all()
        np.testing.assert_allclose(c, c_ref, rtol=RTOL)
        scale = np.abs(g_ref).max(axis=1, keepdims=True)
        assert (np.abs(g - g_ref) / scale).max() < RTOL
        print(f"model {model} N={N}: bit-identical cost {np.mean(c == c_ref):.3f}, gradient {np.mean(g == g_ref):.3f}")


def test_value_grad_static_rejects_what_it_cannot_hold():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    import mpcb200 as mp
    from mpcb200 import _ffi
    N = 12
    prob = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_TABLE, n_roads=2)
    t = _inputs(mp, torch, 8, N, 5)
    c = torch.empty(8, dtype=torch.float64, device=prob.device)
    g = torch.empty(8, 2 * N, dtype=torch.float64, device=prob.device)

    # B beyond the stated maximum: refused before anything is launched (the small tensors are never read)
    import ctypes as C
    from mpcb200.solver import _dev_ptr
    f64 = torch.float64
    with pytest.raises(mp.MpcbError, match="MPCB_STATIC_MAX_B"):
        _ffi.check(prob._lib.mpcb_value_grad_static(
            prob._h, _ffi.STATIC_MAX_B + 1, _dev_ptr(t["x0"], f64, "x0"), _dev_ptr(t["U0"], f64, "U"),
            _dev_ptr(prob.p, f64, "p"), _dev_ptr(t["road"], f64, "road"),
            _dev_ptr(t["road_idx"], torch.int32, "road_idx"), None, None, None, _dev_ptr(c, f64, "c"),
            _dev_ptr(g, f64, "g"), C.c_void_p(0)), prob._h, "mpcb_value_grad_static")
    # layouts that need allocated scratch are refused, not silently served by another path
    per = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_PER_INSTANCE)
    road_pi = t["road"][t["road_idx"].long()].contiguous()
    with pytest.raises(mp.MpcbError, match="MPCB_ROAD_SHARED"):
        per.value_grad_static(t["x0"], t["U0"], road_pi, c, g)
    # the existing entry point still serves them
    per.value_grad(t["x0"], t["U0"], road_pi)
