import sys, warnings
sys.path.insert(0, ".")
import torch
import botorch_b200 as bb
from botorch_b200.synthetic import make_problem
from oracle.qei import OracleQLogEI
from tests.util import oracle_gp, candidates, rel_err
from tests.test_gpu_api import _model
from tests.test_gpu_ei_family import _incumbent
cuda = torch.device("cuda:0")
for cfg, kw, q, B in [("C1", {"noise": 1e-4}, 2, 6), ("C2", {"noise": 1e-4}, 4, 8), ("C2", {"n": 400, "noise": 1e-4}, 3, 6), ("C3", {"n": 300, "noise": 1e-4}, 8, 6), ("C1", {"n": 150, "noise": 1e-4}, 3, 6)]:
    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    X = candidates(p, B, q=q)
    bf = _incumbent(ogp, X)
    o = OracleQLogEI(ogp, bf, num_samples=p.mc_samples, seed=1234)
    acqf = bb.qLogExpectedImprovement(_model(p, cuda), bf.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234))
    Xd = X.to(cuda).requires_grad_(True)
    v = acqf(Xd); (g,) = torch.autograd.grad(v.sum(), Xd)
    a_ref, g_ref = o.value_and_grad(X)
    print(cfg, kw, "val", rel_err(v, a_ref, scale=0.0), "grad", rel_err(g, g_ref))
