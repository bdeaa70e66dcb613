"""Run the sequential-greedy optimisation of tests/test_gpu_api.py twice in one process and print a digest of the candidates:
identical digests within and across processes mean the optimiser path is deterministic.  Usage: python tools/determinism_check.py"""
import hashlib
import sys

import torch

sys.path.insert(0, ".")
import botorch_b200 as bb  # noqa: E402
from botorch_b200.synthetic import make_problem  # noqa: E402
from tests.test_gpu_api import _acqf, _model  # noqa: E402

cuda = torch.device("cuda:0")
p = make_problem("C2", n=50)
for rep in range(2):
    torch.manual_seed(0)  # initialize_q_batch draws its multinomial selection from torch's global CPU RNG, as the reference does
    model = _model(p, cuda)
    acqf = _acqf(p, model, cuda)
    bounds = torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double().to(cuda)
    cands, vals = bb.optimize_acqf(acqf, bounds, q=3, num_restarts=4, raw_samples=64, sequential=True,
                                   options={"maxiter": 20, "seed": 0})  # fmt: skip
    print(rep, hashlib.sha256(cands.cpu().numpy().tobytes()).hexdigest()[:16], [f"{v:.15f}" for v in vals.tolist()])
