"""Time the one-time factor setup (bo_gp_setup, bo_nei_setup) at a few training-set sizes."""
import sys
import time

import torch

sys.path.insert(0, ".")
from botorch_b200 import engine  # noqa: E402
from botorch_b200.synthetic import make_problem  # noqa: E402
from tests.util import gp_state, oracle_gp  # noqa: E402,F401

dev = torch.device("cuda:0")
for n in [int(a) for a in sys.argv[1:]] or [2000, 8000]:
    p = make_problem("C3", n=n)
    h = p.hypers
    y = (p.train_Y - p.train_Y.mean()) / p.train_Y.std()
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        st = engine.GPState(p.train_X.to(dev), y.to(dev), p.kernel, h["lengthscale"].to(dev), h["noise"].to(dev), h["mean"].to(dev))
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        Zb = torch.randn(256, n, dtype=torch.float64, device=dev)
        nei = engine.NEIState(st, p.train_X.to(dev), Zb)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        print(f"n={n} rep={rep} gp_setup {1e3*(t1-t0):.1f} ms ({n**3/3*2/(t1-t0)/1e12:.2f} TF on 2n^3/3)  nei_setup {1e3*(t2-t1):.1f} ms", flush=True)
        del st, nei
