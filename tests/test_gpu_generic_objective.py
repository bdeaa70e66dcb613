"""GPU parity of the sample-path acquisition classes that VERDICT r1 listed as unsupported: a non-linear
``GenericMCObjective`` (acquisition/objective.py:344-384) and outcome constraints on the fixed-incumbent
``qLogExpectedImprovement`` / ``qExpectedImprovement`` (acquisition/logei.py:138-234, monte_carlo.py:305-414), against
oracle/constrained.py + torch autograd.  Values 1e-9, gradients 1e-9 of their largest entry."""
import warnings

import pytest
import torch

from oracle.constrained import OracleConstrainedQLogEI, OracleConstrainedQLogNEI
from oracle.gp import GPHyper, OracleGP
from tests.test_gpu_constrained import _check
from tests.test_gpu_constrained import _data as _raw_data

pytestmark = pytest.mark.gpu


def _data(n, d, seed):
    """tests/test_gpu_constrained._data with observation noise 0.1: at noise 0.02 the ORACLE's own gradient moves by up to
    2e-9 (relative) under a 1e-14 perturbation of X at these sizes -- most candidates sit 40 nats into the log tail -- so the
    1e-9 bar would test rounding noise; at 0.1 and d = 4 that floor is ~1e-13 (measured)."""
    X, ys, hy = _raw_data(n, d, seed)
    for h in hy:
        h["noise"] = 0.1
    return X, ys, hy


def _models(cuda, X, ys, hy):
    import botorch_b200 as bb

    gms = [bb.SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["ls"], noise=h["noise"], mean_constant=h["mean"])
           for y, h in zip(ys, hy)]  # fmt: skip
    ogps = [OracleGP(X, y.unsqueeze(-1), GPHyper("rbf", h["ls"], torch.tensor(h["noise"], dtype=torch.float64),
                                                 torch.tensor(h["mean"], dtype=torch.float64), None), standardize=True)
            for y, h in zip(ys, hy)]  # fmt: skip
    return gms, ogps


def _quadratic(Y, X=None):  # non-linear in the outcome, and reads the inputs
    return -((Y[..., 0] - 0.3) ** 2) + 0.1 * X[..., 0]


def _two_outcomes(Y, X=None):
    return Y[..., 0] - 0.5 * torch.tanh(Y[..., 1]) ** 2


@pytest.mark.parametrize("q,B,S", [(1, 4, 32), (3, 6, 64)])
def test_generic_objective_single_model_qlognei(cuda, q, B, S):
    import botorch_b200 as bb
    from botorch_b200.constrained import qLogNEIModelList

    X, ys, hy = _data(40, 4, seed=1)
    gms, ogps = _models(cuda, X, ys[:1], hy[:1])
    acqf = bb.qLogNoisyExpectedImprovement(gms[0], X.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=7),
                                           objective=bb.GenericMCObjective(_quadratic))  # fmt: skip
    assert isinstance(acqf, qLogNEIModelList)
    oracle = OracleConstrainedQLogNEI(ogps, X, [], 1e-3, num_samples=S, seed=7, objective=_quadratic)
    g = torch.Generator().manual_seed(2)
    _check(cuda, acqf, oracle, torch.rand(B, q, 4, generator=g, dtype=torch.float64))


def test_generic_objective_over_two_outcomes_with_a_constraint(cuda):
    import botorch_b200 as bb

    X, ys, hy = _data(60, 4, seed=3)
    gms, ogps = _models(cuda, X, ys, hy)
    cons = [lambda Z: Z[..., 2] - 0.4]
    S = 64
    acqf = bb.qLogNoisyExpectedImprovement(bb.ModelListGP(*gms), X.to(cuda),
                                           sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=11),
                                           objective=bb.GenericMCObjective(_two_outcomes), constraints=cons, eta=0.05)  # fmt: skip
    oracle = OracleConstrainedQLogNEI(ogps, X, cons, 0.05, num_samples=S, seed=11, objective=_two_outcomes)
    g = torch.Generator().manual_seed(4)
    _check(cuda, acqf, oracle, torch.rand(5, 2, 4, generator=g, dtype=torch.float64))
    # the plain counterpart (qNEI): same samples, relu x sigmoid, amax, mean
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        plain = bb.qNoisyExpectedImprovement(bb.ModelListGP(*gms), X.to(cuda),
                                             sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=11),
                                             objective=bb.GenericMCObjective(_two_outcomes), constraints=cons, eta=0.05,
                                             prune_baseline=False)  # fmt: skip
    oplain = OracleConstrainedQLogNEI(ogps, X, cons, 0.05, num_samples=S, seed=11, objective=_two_outcomes, plain=True,
                                      fat=False)  # fmt: skip
    _check(cuda, plain, oplain, torch.rand(5, 2, 4, generator=g, dtype=torch.float64))


@pytest.mark.parametrize("fat", [True, False])
def test_constrained_qlogei_matches_oracle(cuda, fat):
    """Fixed incumbent, two outcomes: objective = outcome 0, outcome 1 constrained."""
    import botorch_b200 as bb
    from botorch_b200.constrained import qLogEIModelList

    X, ys, hy = _data(60, 4, seed=5)
    gms, ogps = _models(cuda, X, ys[:2], hy[:2])
    cons = [lambda Z: Z[..., 1] - 0.1]
    S, best_f = 64, float(ys[0].max())
    w = torch.tensor([1.0, 0.0], dtype=torch.float64)
    acqf = bb.qLogExpectedImprovement(bb.ModelListGP(*gms), best_f, sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=3),
                                      objective=bb.LinearMCObjective(w.to(cuda)), constraints=cons, eta=0.05, fat=fat)  # fmt: skip
    assert isinstance(acqf, qLogEIModelList)
    oracle = OracleConstrainedQLogEI(ogps, best_f, cons, 0.05, num_samples=S, seed=3, fat=fat)
    g = torch.Generator().manual_seed(6)
    _check(cuda, acqf, oracle, torch.rand(6, 3, 4, generator=g, dtype=torch.float64))


def test_constraint_on_the_single_outcome_of_qlogei_and_qei(cuda):
    """One model, the constraint reads the same outcome as the objective (m = 1 samples: S x B x q x 1)."""
    import botorch_b200 as bb

    X, ys, hy = _data(40, 4, seed=8)
    gms, ogps = _models(cuda, X, ys[:1], hy[:1])
    cons = [lambda Z: Z[..., 0] - 1.2]  # feasible while the outcome stays below 1.2
    S, best_f = 32, float(ys[0].median())
    g = torch.Generator().manual_seed(9)
    Xc = torch.rand(4, 2, 4, generator=g, dtype=torch.float64)
    acqf = bb.qLogExpectedImprovement(gms[0], best_f, sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=5),
                                      constraints=cons, eta=0.1)  # fmt: skip
    _check(cuda, acqf, OracleConstrainedQLogEI(ogps, best_f, cons, 0.1, num_samples=S, seed=5), Xc)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        plain = bb.qExpectedImprovement(gms[0], best_f, sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=5),
                                        constraints=cons, eta=0.1)  # fmt: skip
    _check(cuda, plain, OracleConstrainedQLogEI(ogps, best_f, cons, 0.1, num_samples=S, seed=5, plain=True, fat=False), Xc)
    # without constraints and with an identity objective the sample path equals the all-in-kernel qLogEI
    fused = bb.qLogExpectedImprovement(gms[0], best_f, sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=5))
    generic = bb.qLogExpectedImprovement(gms[0], best_f, sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=5),
                                         objective=bb.GenericMCObjective(lambda Y, X=None: Y[..., 0]))  # fmt: skip
    a, b = fused(Xc.to(cuda)), generic(Xc.to(cuda))
    assert float((a - b).abs().max()) < 1e-10 * float(a.abs().max())


def test_generic_objective_under_optimize_acqf(cuda):
    import botorch_b200 as bb

    X, ys, hy = _data(40, 4, seed=10)
    gms, _ = _models(cuda, X, ys[:1], hy[:1])
    acqf = bb.qLogNoisyExpectedImprovement(gms[0], X.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([32]), seed=1),
                                           objective=bb.GenericMCObjective(_quadratic))  # fmt: skip
    bounds = torch.stack([torch.zeros(4), torch.ones(4)]).double().to(cuda)
    cand, val = bb.optimize_acqf(acqf, bounds, q=2, num_restarts=4, raw_samples=32, options={"maxiter": 10, "seed": 0})
    assert cand.shape == (2, 4) and bool(torch.isfinite(val).all())
    assert bool((cand >= 0).all()) and bool((cand <= 1).all())
    with torch.no_grad():
        assert float(acqf(cand.unsqueeze(0))) >= float(acqf(torch.rand(64, 2, 4, dtype=torch.float64, device=cuda)).median())
