"""Summarise an MPCB_ROUND_LOG file (round, gradient requests, psi requests, T ms, E ms per line): bulk vs tail."""
import sys

import numpy as np


def main(path, thr=40000):
    a = np.loadtxt(path)
    if a.ndim == 1:
        a = a[None]
    req = a[:, 1] + a[:, 2]
    bulk = req > thr
    out = dict(rounds=int(len(a)), bulk_rounds=int(bulk.sum()), bulk_T_ms=float(a[bulk, 3].sum()), bulk_E_ms=float(a[bulk, 4].sum()),
               tail_T_ms=float(a[~bulk, 3].sum()), tail_E_ms=float(a[~bulk, 4].sum()),
               bulk_T_per_round=float(a[bulk, 3].mean()) if bulk.any() else 0.0,
               bulk_E_per_round=float(a[bulk, 4].mean()) if bulk.any() else 0.0,
               tail_T_per_round=float(a[~bulk, 3].mean()) if (~bulk).any() else 0.0,
               tail_E_per_round=float(a[~bulk, 4].mean()) if (~bulk).any() else 0.0,
               requests_total=float(req.sum()), bulk_requests=float(req[bulk].sum()))
    print(path, {k: (round(v, 3) if isinstance(v, float) else v) for k, v in out.items()})
    # a few sample rounds
    for r in (5, 20, 50, 100, 150, 200, 300, 400, 600, 1000, 2000):
        if r < len(a):
            print(f"   round {r}: grad {int(a[r, 1])} psi {int(a[r, 2])} T {a[r, 3]:.3f} ms E {a[r, 4]:.3f} ms")


if __name__ == "__main__":
    for p in sys.argv[1:]:
        main(p)
