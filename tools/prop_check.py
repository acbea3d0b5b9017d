import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch, mpcb200
from mpcb200 import workloads as wl
N, B = 20, 4096
w = wl.tracking_batch(B, N, 100, seed=2)
prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
U, _, st = prob.solve(w["x0"], w["road"], w["U0"].copy(), road_idx=w["road_idx"])
_, _, c0 = prob.rollout(w["x0"], w["U0"], w["road"], road_idx=w["road_idx"])
_, _, c1 = prob.rollout(w["x0"], U, w["road"], road_idx=w["road_idx"])
c0, c1, psi = c0.cpu().numpy(), c1.cpu().numpy(), st.psi.cpu().numpy()
bad = np.where(c1 > c0 + 1e-12)[0]
print("violators", bad, "c0", c0[bad], "c1", c1[bad], "psi", psi[bad], "status", st.status.cpu().numpy()[bad],
      "iters", st.inner_iterations.cpu().numpy()[bad], "outer", st.outer_iterations.cpu().numpy()[bad])
print("psi vs rollout max rel", np.abs(c1 - psi).max())
np.save("gpurun_out/prop_U.npy", U.cpu().numpy())
