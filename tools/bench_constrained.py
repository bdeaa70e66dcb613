"""Constrained qLogNEI at the C3 shape: two outcomes (objective + one constrained outcome), n=2000, d=20,
q=8, S=256, 512 restarts; forward + backward through the public API, device-timed; the CPU oracle on a bounded sample.
Usage: python tools/bench_constrained.py [--no-cpu]"""
import json
import sys
import time

import torch

sys.path.insert(0, ".")
import botorch_b200 as bb  # noqa: E402
from botorch_b200.synthetic import make_problem  # noqa: E402

dev = torch.device("cuda:0")
p = make_problem("C3")
h = p.hypers
g = torch.Generator().manual_seed(0)
X = p.train_X
y_con = (X[:, :4].sum(-1) - 2.0 + 0.05 * torch.randn(X.shape[0], generator=g, dtype=torch.float64)).unsqueeze(-1)
cons = [lambda Z: Z[..., 1] - 0.25]
S, B, q, d = p.mc_samples, p.num_restarts, p.q, p.d


def build():
    gps = [bb.SingleTaskGP(X.to(dev), y.to(dev), lengthscale=h["lengthscale"], noise=h["noise"], mean_constant=h["mean"])
           for y in (p.train_Y, y_con)]  # fmt: skip
    return bb.qLogNoisyExpectedImprovement(
        bb.ModelListGP(*gps), X.to(dev), sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=1234),
        objective=bb.LinearMCObjective(torch.tensor([1.0, 0.0], dtype=torch.float64, device=dev)), constraints=cons, eta=1e-2)  # fmt: skip


torch.cuda.synchronize()
t0 = time.perf_counter()
acqf = build()
torch.cuda.synchronize()
setup_s = time.perf_counter() - t0
Xc = torch.rand(B, q, d, dtype=torch.float64, generator=g).to(dev)


def step():
    Xg = Xc.detach().requires_grad_(True)
    v = acqf(Xg)
    (gr,) = torch.autograd.grad(v.sum(), Xg)
    return v, gr


for _ in range(5):
    step()
K = 20
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
w0 = time.perf_counter()
e0.record()
for _ in range(K):
    v, gr = step()
e1.record()
torch.cuda.synchronize()
wall = (time.perf_counter() - w0) / K
ms = e0.elapsed_time(e1) / K
with torch.no_grad():
    for _ in range(3):
        acqf(Xc)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        acqf(Xc)
    e1.record()
    torch.cuda.synchronize()
ms_fwd = e0.elapsed_time(e1) / K
out = {"config": f"C3 shape, 2 outcomes + 1 constraint: n={X.shape[0]} d={d} q={q} S={S} restarts={B}", "setup_s": setup_s,
       "fwd_bwd_ms": ms, "fwd_bwd_wall_ms": wall * 1e3, "fwd_only_ms": ms_fwd, "candidate_evals_per_s": B / (ms * 1e-3),
       "finite": bool(torch.isfinite(v).all() and torch.isfinite(gr).all()),
       "acq_mean": float(v.detach().mean())}  # fmt: skip
if "--no-cpu" not in sys.argv:
    from oracle.constrained import OracleConstrainedQLogNEI
    from oracle.gp import GPHyper, OracleGP

    hy = GPHyper("rbf", h["lengthscale"], torch.as_tensor(h["noise"], dtype=torch.float64),
                 torch.as_tensor(h["mean"], dtype=torch.float64), None)  # fmt: skip
    t0 = time.perf_counter()
    ogps = [OracleGP(X, y, hy, standardize=True) for y in (p.train_Y, y_con)]
    orc = OracleConstrainedQLogNEI(ogps, X, cons, 1e-2, obj_idx=0, obj_weight=1.0, num_samples=S, seed=1234)
    t1 = time.perf_counter()
    nb = 4
    ov, og = orc.value_and_grad(Xc[:nb].cpu())
    t2 = time.perf_counter()
    out["cpu_oracle"] = {"setup_s": t1 - t0, "restarts": nb, "seconds": t2 - t1, "candidate_evals_per_s": nb / (t2 - t1),
                         "cores": torch.get_num_threads(),
                         "value_rel_err": float((v[:nb].cpu() - ov).abs().max()),
                         "grad_rel_err": float((gr[:nb].cpu() - og).abs().max() / og.abs().max())}  # fmt: skip
print(json.dumps(out))
