"""How often does the q x q Cholesky of the cached update need jitter (info > 0) or fail (info < 0) along a real
optimize_acqf run?  (the jitter branch makes the objective discontinuous where greedy candidates land.)
Usage: python tools/jitter_events.py [C2 C3]"""
import json
import sys
import warnings

import torch

sys.path.insert(0, ".")
import botorch_b200 as bb  # noqa: E402
from botorch_b200.synthetic import make_problem  # noqa: E402

dev = torch.device("cuda:0")
out = {}
for cfg in sys.argv[1:] or ["C2", "C3"]:
    p = make_problem(cfg)
    h = p.hypers
    model = bb.SingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev), lengthscale=h["lengthscale"], noise=h["noise"],
                            mean_constant=h["mean"])  # fmt: skip
    acqf = bb.qLogNoisyExpectedImprovement(model, p.train_X.to(dev),
                                           sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234))  # fmt: skip
    acqf.strict_numerics(True)  # look at the flags of every call
    rec = {"calls": 0, "tbatch_evals": 0, "calls_with_jitter": 0, "tbatches_with_jitter": 0, "max_retries": 0,
           "calls_failed": 0, "tbatches_failed": 0}
    orig = acqf._failed_tbatches

    def spy(info, _orig=orig, _rec=rec, _acqf=acqf):
        full = info if info.numel() != 2 or info.dim() != 1 else _acqf._nei.last_graph_info
        _rec["calls"] += 1
        _rec["tbatch_evals"] += int(full.shape[-1])
        nj, nf = int((full > 0).any(0).sum()), int((full < 0).any(0).sum())
        _rec["calls_with_jitter"] += nj > 0
        _rec["tbatches_with_jitter"] += nj
        _rec["max_retries"] = max(_rec["max_retries"], int(full.max()))
        _rec["calls_failed"] += nf > 0
        _rec["tbatches_failed"] += nf
        return _orig(info)

    acqf._failed_tbatches = spy
    bounds = torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double().to(dev)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for sequential in (False, True):
            cand, val = bb.optimize_acqf(acqf, bounds, q=p.q, num_restarts=p.num_restarts if cfg != "C3" else 64,
                                         raw_samples=p.raw_samples if cfg != "C3" else 4096,
                                         options={"seed": 0, "maxiter": 50}, sequential=sequential)  # fmt: skip
            rec[f"best_value_{'sequential' if sequential else 'joint'}"] = float(val.sum())
    out[cfg] = rec
print(json.dumps(out))
