import sys, warnings
sys.path.insert(0, ".")
import torch
from tests.test_gpu_edges import _problem
from tests.util import oracle_gp, oracle_acq, gp_state
from oracle import qmc, safe_math
from botorch_b200.engine import NEIState
cuda = torch.device("cuda:0")
n, d, q, S, B, kernel, M = 129, 3, 9, 31, 5, "matern52", 2
p = _problem(n, d, q, S, kernel, M)
ogp = oracle_gp(p); oacq = oracle_acq(p, ogp)
st = gp_state(p, ogp, cuda)
nei = NEIState(st, p.train_X.to(cuda), oacq.Z_b.to(cuda))
bounds = torch.stack([torch.zeros(d), torch.ones(d)]).double()
X = qmc.sobol_box(bounds, B, q, seed=77)
Zq = oacq.Z_q(q).to(cuda).contiguous()
acq, info, work = nei.forward(X.to(cuda), Zq, oacq.tau_relu, oacq.tau_max, with_grad=True)
print("info", info.cpu())
with warnings.catch_warnings(record=True) as ws:
    warnings.simplefilter("always")
    obj = oacq.samples_faithful(X)  # S x B x M x q
    print("oracle warnings", [str(w.message)[:80] for w in ws])
li = safe_math.log_improvement(obj, oacq.best_f.unsqueeze(1), tau=oacq.tau_relu)
per = safe_math.logmeanexp(safe_math.fatmax(li, dim=-1, tau=oacq.tau_max), dim=0)  # B x M
print("gpu ", acq.cpu())
print("ref ", per.mT)
print("diff", (acq.cpu() - per.mT).abs())
print("best_f diff", (nei.best_f.cpu() - oacq.best_f.mT).abs().max())
print("baseline_L diff", (nei.baseline_L.cpu() - oacq.baseline_L).abs().max())
