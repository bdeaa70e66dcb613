"""One profiled invocation of (a) the constrained two-outcome qLogNEI step at the C3 shape and (b) the SAAS potential at the
C4 shape with 64 chains, bracketed by cudaProfilerStart/Stop.  Run under
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/ncu_new_paths.py
Usage: python tools/ncu_new_paths.py [constrained|saas]"""
import sys

import torch

sys.path.insert(0, ".")
import botorch_b200 as bb  # noqa: E402
from botorch_b200.fully_bayesian import SaasPotential  # noqa: E402
from botorch_b200.synthetic import make_problem  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "constrained"
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
rt = torch.cuda.cudart()
if which == "constrained":
    p = make_problem("C3")
    h = p.hypers
    X = p.train_X
    y_con = (X[:, :4].sum(-1) - 2.0 + 0.05 * torch.randn(X.shape[0], generator=g, dtype=torch.float64)).unsqueeze(-1)
    gps = [bb.SingleTaskGP(X.to(dev), y.to(dev), lengthscale=h["lengthscale"], noise=h["noise"], mean_constant=h["mean"])
           for y in (p.train_Y, y_con)]  # fmt: skip
    acqf = bb.qLogNoisyExpectedImprovement(
        bb.ModelListGP(*gps), X.to(dev), sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234),
        objective=bb.LinearMCObjective(torch.tensor([1.0, 0.0], dtype=torch.float64, device=dev)),
        constraints=[lambda Z: Z[..., 1] - 0.25], eta=1e-2)  # fmt: skip
    Xc = torch.rand(p.num_restarts, p.q, p.d, dtype=torch.float64, generator=g).to(dev)

    def step():
        Xg = Xc.detach().requires_grad_(True)
        torch.autograd.grad(acqf(Xg).sum(), Xg)
else:
    n, d, C = 100, 30, 64
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    y = torch.sin(2 * torch.pi * X[:, 0]) + 0.1 * torch.randn(n, generator=g, dtype=torch.float64)
    pot = SaasPotential(X.to(dev), ((y - y.mean()) / y.std()).to(dev), num_chains=C)
    U = (torch.rand(C, d + 4, generator=g, dtype=torch.float64) - 0.5).to(dev)

    def step():
        pot(U)
for _ in range(4):
    step()
torch.cuda.synchronize()
rt.cudaProfilerStart()
step()
torch.cuda.synchronize()
rt.cudaProfilerStop()
print("ok", which)
