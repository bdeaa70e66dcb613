import sys, warnings
sys.path.insert(0, ".")
import torch
from botorch_b200.synthetic import make_problem
from tests.util import oracle_gp, oracle_acq, gp_state, candidates, rel_err
from tests.test_gpu_qlognei import _grad_rel
from botorch_b200.engine import NEIState
cuda = torch.device("cuda:0")
for cfg, kw, q, B in [("C1", {"noise": 1e-4}, 2, 6), ("C2", {"noise": 1e-4}, 4, 8), ("C3", {"n": 300, "noise": 1e-4}, 8, 6), ("C2", {"n": 400, "noise": 1e-4}, 3, 6)]:
    p = make_problem(cfg, **kw)
    with warnings.catch_warnings(record=True) as ws:
        warnings.simplefilter("always")
        ogp = oracle_gp(p); oacq = oracle_acq(p, ogp)
        st = gp_state(p, ogp, cuda)
        nei = NEIState(st, p.train_X.to(cuda), oacq.Z_b.to(cuda))
        X = candidates(p, B, q=q)
        Zq = oacq.Z_q(q).to(cuda).contiguous()
        acq, info, work = nei.forward(X.to(cuda), Zq, 1e-6, 1e-2, with_grad=True)
        g = nei.backward(X.to(cuda), Zq, 1e-6, 1e-2, torch.ones_like(acq), work)[0]
        a_ref, g_ref = oacq.value_and_grad(X, mode="faithful")
        a_unc = oacq.forward(X, mode="uncached")
        print(cfg, kw, "jitter gp", st.jitter.tolist(), "nei", nei.jitter.tolist(), "info", info.cpu().tolist())
        print("   val rel err", rel_err(acq[0], a_ref, scale=0.0), "grad", _grad_rel(g, g_ref), "oracle faithful-vs-uncached", rel_err(a_unc, a_ref, scale=0.0))
        print("   warnings", sorted({str(w.message)[:60] for w in ws}))
