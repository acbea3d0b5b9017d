import os, sys, json, time
sys.path.insert(0, os.getcwd())
import torch, numpy as np
import mpcb200
from mpcb200 import _ffi
import ctypes
_L = ctypes.CDLL(_ffi.LIB_PATH)
_ffi.SIGNATURES = {k: v for k, v in _ffi.SIGNATURES.items() if hasattr(_L, k)}  # older builds of the ABI (A/B via MPCB_LIB)
from mpcb200 import scenarios as sc, workloads as wl
B, N = 8192, 40
lc = wl.lane_change_batch(B, N, 100, seed=3, n_obs=2)
prob = mpcb200.BatchedMpc(N=N, S=100, model_id=1, road_mode=mpcb200.ROAD_BEZIER, flags=mpcb200.FLAG_LANE | mpcb200.FLAG_OBST, n_obs=2, v_ref=20.0)
Uc, st_c, x1, Uw, st_w = sc.event_triggered_resolve(prob, lc["x0"], lc["ctrl"], lc["U0"], obs=lc["obs"], seed=3)
dev = prob.device
ctrl, obs = torch.tensor(lc["ctrl"], device=dev), torch.tensor(lc["obs"], device=dev)
for rep in range(2):
    U = Uc.clone(); torch.cuda.synchronize(); t = time.perf_counter()
    _, _, st = prob.solve(x1, ctrl, U, obs=obs); torch.cuda.synchronize(); dt = time.perf_counter() - t
print(json.dumps({"lib": os.path.basename(_ffi.LIB_PATH), "pit": os.environ.get("MPCB_PIT", "1"), "loop": os.environ.get("MPCB_LOOP", "host"), "ms": dt * 1e3, "profile": prob.last_solve_profile(), "inner_mean": float(st.inner_iterations.double().mean().item()), "conv": float((st.status == 1).double().mean().item())}))
