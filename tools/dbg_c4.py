import sys, warnings, torch
sys.path.insert(0, '.')
import botorch_b200 as bb
from botorch_b200.synthetic import make_problem
from tests.util import candidates, oracle_acq, oracle_gp, rel_err
warnings.simplefilter("always")
dev = torch.device("cuda:0")
for M, B in ((5, 7), (6, 9), (5, 9), (6, 7)):
    p = make_problem("C4", M=M)
    ogp = oracle_gp(p); oacq = oracle_acq(p, ogp)
    m = bb.SaasFullyBayesianSingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev)); m.load_mcmc_samples(p.hypers)
    sampler = bb.SobolQMCNormalSampler(sample_shape=torch.Size([p.mc_samples]), seed=1234)
    acqf = bb.qLogNoisyExpectedImprovement(m, p.train_X.to(dev), sampler=sampler)
    X = candidates(p, B, q=4)
    nei = acqf._nei
    acqf._set_sampler(4)
    acq, info, _ = nei.forward(X.to(dev), acqf._Zq, 1e-6, 1e-2, with_grad=False)
    # per-model oracle values
    obj = oacq.samples_faithful(X)
    from oracle import safe_math
    per = safe_math.qlognei_reduction(obj, oacq.best_f.unsqueeze(1), 1e-6, 1e-2)  # B x M
    print(M, B, "gp jitter", m.gp_state.jitter.tolist(), "nei jitter", nei.jitter.tolist(), "info", info.unique().tolist())
    print("   per-model rel err", ((acq.cpu().mT - per).abs() / per.abs()).max(dim=0).values.tolist())
    print("   L err", rel_err(m.gp_state.L, ogp.caches()[0]), "Lb err", rel_err(nei.baseline_L, oacq.baseline_L), "bestf", rel_err(nei.best_f.mT, oacq.best_f),
          "Zq", float((acqf._Zq.cpu() - oacq.Z_q(4)).abs().max()), "Zb", float((nei.Z_b.cpu() - oacq.Z_b).abs().max()))
