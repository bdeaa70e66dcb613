"""GPU parity of qLogNEI with outcome constraints over a ModelListGP (bo_mc_reduce_fwd / _bwd, external
sample adjoints through bo_qlognei_bwd) against the oracle restatement (oracle/constrained.py + torch autograd)."""
import pytest
import torch

from oracle.constrained import OracleConstrainedQLogNEI
from oracle.gp import GPHyper, OracleGP
from tests.util import rel_err

pytestmark = pytest.mark.gpu

# Problems are d >= 3: with d = 2 and these lengthscales the baseline posterior covariance is close to singular and the
# ORACLE's own value/gradient move by 1e-10..1e-9 under a 1e-14 perturbation of X (measured), i.e. the comparison would
# test rounding noise; at d >= 3 that noise floor is ~1e-12 and the 1e-9 bar is meaningful.
TOL = 1e-9


def _data(n, d, seed):
    g = torch.Generator().manual_seed(seed)
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    y0 = torch.sin(4.0 * X[:, 0]) + X[:, 1] ** 2 + 0.05 * torch.randn(n, generator=g, dtype=torch.float64)
    y1 = X[:, 0] - 0.8 * X[:, 1] + 0.05 * torch.randn(n, generator=g, dtype=torch.float64)  # feasible where <= threshold
    y2 = torch.cos(3.0 * X[:, -1]) + 0.05 * torch.randn(n, generator=g, dtype=torch.float64)
    hy = [
        {"ls": 0.3 + torch.rand(d, generator=g, dtype=torch.float64), "noise": 0.02 + 0.01 * k, "mean": 0.1 * k}
        for k in range(3)
    ]
    return X, [y0, y1, y2], hy


def _build(cuda, X, ys, hy, S, seed, constraints, eta, obj_idx=0, obj_w=1.0, fat=True, X_pending=None):
    import botorch_b200 as bb

    m = len(ys)
    gms = [bb.SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["ls"], noise=h["noise"], mean_constant=h["mean"])
           for y, h in zip(ys, hy)]  # fmt: skip
    w = torch.zeros(m, dtype=torch.float64)
    w[obj_idx] = obj_w
    acqf = bb.qLogNoisyExpectedImprovement(
        bb.ModelListGP(*gms), X.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([S]), seed=seed),
        objective=bb.LinearMCObjective(w.to(cuda)), constraints=constraints, eta=eta, fat=fat,
        X_pending=None if X_pending is None else X_pending.to(cuda))  # fmt: skip
    ogps = [OracleGP(X, y.unsqueeze(-1), GPHyper("rbf", h["ls"], torch.tensor(h["noise"], dtype=torch.float64),
                                                 torch.tensor(h["mean"], dtype=torch.float64), None), standardize=True)
            for y, h in zip(ys, hy)]  # fmt: skip
    oracle = OracleConstrainedQLogNEI(ogps, X, constraints or [], eta, obj_idx=obj_idx, obj_weight=obj_w, num_samples=S,
                                      seed=seed, fat=fat)  # fmt: skip
    return acqf, oracle


def _check(cuda, acqf, oracle, Xc, tol=TOL):
    Xg = Xc.to(cuda).requires_grad_(True)
    val = acqf(Xg)
    (grad,) = torch.autograd.grad(val.sum(), Xg)
    oval, ograd = oracle.value_and_grad(Xc)
    assert rel_err(val, oval, scale=1.0) < tol
    assert rel_err(grad, ograd, scale=float(ograd.abs().max())) < tol
    with torch.no_grad():
        assert rel_err(acqf(Xc.to(cuda)), oval, scale=1.0) < tol


@pytest.mark.parametrize("n,d,q,B,S,m", [(30, 3, 1, 5, 32, 2), (60, 3, 3, 7, 64, 2), (120, 4, 4, 16, 128, 3)])
def test_constrained_qlognei_matches_oracle(cuda, n, d, q, B, S, m):
    X, ys, hy = _data(n, d, seed=n)
    ys, hy = ys[:m], hy[:m]
    cons = [lambda Z: Z[..., 1] - 0.6]
    if m == 3:
        cons.append(lambda Z: 0.3 * Z[..., 0] - Z[..., 2] - 1.5)  # couples the objective outcome with the third one
    acqf, oracle = _build(cuda, X, ys, hy, S, seed=7, constraints=cons, eta=torch.tensor([1e-2, 5e-2][: len(cons)]))
    assert type(acqf).__name__ == "qLogNEIModelList"
    assert rel_err(acqf._best_f, oracle.best_f, scale=1.0) < TOL
    Xc = torch.rand(B, q, d, dtype=torch.float64, generator=torch.Generator().manual_seed(3))
    _check(cuda, acqf, oracle, Xc)


def test_constrained_logexpit_minimisation_and_pending_points(cuda):
    X, ys, hy = _data(50, 3, seed=4)
    cons = [lambda Z: Z[..., 1] - 0.7]
    acqf, oracle = _build(cuda, X, ys[:2], hy[:2], 64, seed=3, constraints=cons, eta=0.05, obj_w=-1.0, fat=False)
    Xc = torch.rand(6, 2, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(5))
    _check(cuda, acqf, oracle, Xc)
    # X_pending is concatenated to every t-batch (concatenate_pending_points): q = 2 + 1
    Xp = torch.rand(1, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(6))
    acqf.set_X_pending(Xp.to(cuda))
    val = acqf(Xc.to(cuda))
    oval = oracle.forward(torch.cat([Xc, Xp.expand(6, 1, 3)], dim=-2))
    assert rel_err(val, oval, scale=1.0) < TOL


def test_single_model_list_equals_plain_qlognei(cuda):
    """A one-model ModelListGP without constraints takes the split path (samples -> bo_mc_reduce) and must reproduce the
    fused single-model kernels bit for bit in value and to rounding in the gradient."""
    import botorch_b200 as bb

    X, ys, hy = _data(80, 3, seed=9)
    h = hy[0]
    gm = bb.SingleTaskGP(X.to(cuda), ys[0].unsqueeze(-1).to(cuda), lengthscale=h["ls"], noise=h["noise"], mean_constant=h["mean"])
    smp = lambda: bb.SobolQMCNormalSampler(torch.Size([64]), seed=11)
    plain = bb.qLogNoisyExpectedImprovement(gm, X.to(cuda), sampler=smp())
    split = bb.qLogNoisyExpectedImprovement(bb.ModelListGP(gm), X.to(cuda), sampler=smp(),
                                            objective=bb.LinearMCObjective(torch.ones(1, dtype=torch.float64, device=cuda)))  # fmt: skip
    Xc = torch.rand(9, 3, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(2)).to(cuda)
    Xa, Xb = Xc.clone().requires_grad_(True), Xc.clone().requires_grad_(True)
    va, vb = plain(Xa), split(Xb)
    (ga,) = torch.autograd.grad(va.sum(), Xa)
    (gb,) = torch.autograd.grad(vb.sum(), Xb)
    assert rel_err(vb, va, scale=1.0) < 1e-13
    # the plain route forms d qmax / d f in its forward kernel (bo_nei.local_derivs) and multiplies by the sample weight in
    # the reverse one, the split route forms the same product in one place: one association differs, and this d = 3
    # problem amplifies a last-bit difference of the sample adjoints ~1e4-fold (see the note at the top of the file)
    assert rel_err(gb, ga, scale=float(ga.abs().max())) < 1e-11


def test_constrained_acquisition_under_optimize_acqf(cuda):
    import botorch_b200 as bb

    X, ys, hy = _data(40, 2, seed=21)
    cons = [lambda Z: Z[..., 1] - 0.5]
    acqf, _ = _build(cuda, X, ys[:2], hy[:2], 64, seed=5, constraints=cons, eta=1e-2)
    bounds = torch.stack([torch.zeros(2), torch.ones(2)]).double().to(cuda)
    cand, val = bb.optimize_acqf(acqf, bounds, q=2, num_restarts=4, raw_samples=64, options={"maxiter": 30, "seed": 0})
    assert cand.shape == (2, 2) and bool(torch.isfinite(val))
    with pytest.raises(bb.exceptions.UnsupportedError):
        bb.qLogNoisyExpectedImprovement(acqf.model, X.to(cuda), constraints=cons)  # no single-outcome objective given


def test_multi_output_single_task_gp_is_the_list_of_its_outputs(cuda):
    """SingleTaskGP(train_X, train_Y[n x m]) == m independent single-output models (models/gpytorch.py:405-466): same
    posterior, and the constrained acquisition built on it equals the one built on the explicit list."""
    import botorch_b200 as bb

    X, ys, hy = _data(40, 3, seed=12)
    Y = torch.stack(ys[:2], dim=-1)
    ls = torch.stack([hy[0]["ls"], hy[1]["ls"]])
    noise = torch.tensor([hy[0]["noise"], hy[1]["noise"]], dtype=torch.float64)
    mean = torch.tensor([hy[0]["mean"], hy[1]["mean"]], dtype=torch.float64)
    multi = bb.SingleTaskGP(X.to(cuda), Y.to(cuda), lengthscale=ls, noise=noise, mean_constant=mean)
    assert isinstance(multi, bb.ModelListGP) and multi.num_outputs == 2
    singles = [bb.SingleTaskGP(X.to(cuda), Y[:, k : k + 1].to(cuda), lengthscale=ls[k], noise=float(noise[k]),
                               mean_constant=float(mean[k])) for k in range(2)]  # fmt: skip
    Xt = torch.rand(4, 2, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(1)).to(cuda)
    pm = multi.posterior(Xt)
    assert pm.mean.shape == (4, 2, 2) and pm.variance.shape == (4, 2, 2)
    for k in range(2):
        pk = singles[k].posterior(Xt)
        assert torch.equal(pm.mean[..., k : k + 1], pk.mean) and torch.equal(pm.variance[..., k : k + 1], pk.variance)
    cons = [lambda Z: Z[..., 1] - 0.6]
    kw = dict(objective=bb.LinearMCObjective(torch.tensor([1.0, 0.0], dtype=torch.float64, device=cuda)), constraints=cons, eta=1e-2)
    a = bb.qLogNoisyExpectedImprovement(multi, X.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([32]), seed=2), **kw)
    b = bb.qLogNoisyExpectedImprovement(bb.ModelListGP(*singles), X.to(cuda),
                                        sampler=bb.SobolQMCNormalSampler(torch.Size([32]), seed=2), **kw)  # fmt: skip
    assert torch.equal(a(Xt), b(Xt))
    oracle_models = [OracleGP(X, Y[:, k : k + 1], GPHyper("rbf", ls[k], noise[k], mean[k], None), standardize=True)
                     for k in range(2)]  # fmt: skip
    oracle = OracleConstrainedQLogNEI(oracle_models, X, cons, 1e-2, obj_idx=0, obj_weight=1.0, num_samples=32, seed=2)
    _check(cuda, a, oracle, Xt.cpu())


def test_incumbent_with_and_without_feasible_baseline_points(cuda):
    """test/acquisition/test_logei.py:701-727: with an always-feasible constraint the incumbent is the unconstrained best
    objective; with a never-feasible one it is the 6-sigma lower bound of acquisition/utils.py:137-199 for every sample."""
    import botorch_b200 as bb

    X, ys, hy = _data(30, 3, seed=17)
    feasible_con = lambda Z: -torch.ones_like(Z[..., 0])
    infeasible_con = lambda Z: torch.ones_like(Z[..., 0])
    acqf, oracle = _build(cuda, X, ys[:2], hy[:2], 32, seed=4, constraints=[feasible_con], eta=1e-2)
    assert torch.equal(acqf._best_f, acqf.baseline_obj.amax(dim=-1))
    assert rel_err(acqf._best_f, oracle.best_f, scale=1.0) < TOL

    torch.cuda.manual_seed(123)
    weights = torch.rand(32, 30, dtype=torch.float64, device=cuda).cpu()  # the draw _estimate_objective_lower_bound makes
    torch.cuda.manual_seed(123)
    import botorch_b200.constrained  # noqa: F401

    gms = [bb.SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["ls"], noise=h["noise"], mean_constant=h["mean"])
           for y, h in zip(ys[:2], hy[:2])]  # fmt: skip
    torch.cuda.manual_seed(123)
    acqf = bb.qLogNoisyExpectedImprovement(
        bb.ModelListGP(*gms), X.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([32]), seed=4),
        objective=bb.LinearMCObjective(torch.tensor([1.0, 0.0], dtype=torch.float64, device=cuda)),
        constraints=[infeasible_con], eta=1e-2)  # fmt: skip
    ogps = [OracleGP(X, y.unsqueeze(-1), GPHyper("rbf", h["ls"], torch.tensor(h["noise"], dtype=torch.float64),
                                                 torch.tensor(h["mean"], dtype=torch.float64), None), standardize=True)
            for y, h in zip(ys[:2], hy[:2])]  # fmt: skip
    oracle = OracleConstrainedQLogNEI(ogps, X, [infeasible_con], 1e-2, obj_idx=0, obj_weight=1.0, num_samples=32, seed=4,
                                      convex_weights=weights)  # fmt: skip
    assert bool((acqf._best_f == acqf._best_f[0]).all()) and float(acqf._best_f[0]) <= 0.0
    assert rel_err(acqf._best_f, oracle.best_f, scale=1.0) < TOL
    Xc = torch.rand(4, 2, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(3))
    _check(cuda, acqf, oracle, Xc)


@pytest.mark.parametrize("q,B,S", [(12, 1, 7), (32, 2, 8), (1, 1, 1)])
def test_constrained_edge_shapes(cuda, q, B, S):
    """Wide q-batches (the QT = 16 / 32 instantiations of the reduction kernels), odd sample counts, a single sample, and
    an input that is neither contiguous nor requires grad."""
    X, ys, hy = _data(40, 3, seed=5)
    cons = [lambda Z: Z[..., 1] - 0.9]
    acqf, oracle = _build(cuda, X, ys[:2], hy[:2], S, seed=2, constraints=cons, eta=0.1)
    Xc = torch.rand(B, q, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(q))
    _check(cuda, acqf, oracle, Xc)
    wide = torch.rand(B, q, 6, dtype=torch.float64, generator=torch.Generator().manual_seed(1)).to(cuda)
    view = wide[..., ::2]  # non-contiguous
    with torch.no_grad():
        assert rel_err(acqf(view), oracle.forward(view.cpu().contiguous()), scale=1.0) < TOL


def test_prune_baseline_over_several_outcomes(cuda):
    """prune_inferior_points with constraints over a model list: the kept index set equals the oracle's (same Sobol draw over
    the points x outcomes block), and `prune_baseline=True` builds the acquisition on the pruned set."""
    import botorch_b200 as bb
    from botorch_b200.constrained import prune_inferior_points_list
    from oracle.constrained import prune_inferior_points as oracle_prune

    X, ys, hy = _data(60, 3, seed=23)
    cons = [lambda Z: Z[..., 1] - 0.3]
    gms = [bb.SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["ls"], noise=h["noise"], mean_constant=h["mean"])
           for y, h in zip(ys[:2], hy[:2])]  # fmt: skip
    ogps = [OracleGP(X, y.unsqueeze(-1), GPHyper("rbf", h["ls"], torch.tensor(h["noise"], dtype=torch.float64),
                                                 torch.tensor(h["mean"], dtype=torch.float64), None), standardize=True)
            for y, h in zip(ys[:2], hy[:2])]  # fmt: skip
    model = bb.ModelListGP(*gms)
    obj = bb.LinearMCObjective(torch.tensor([1.0, 0.0], dtype=torch.float64, device=cuda))
    for max_frac in (1.0, 0.05):
        kept = prune_inferior_points_list(model, X.to(cuda), obj, constraints=cons, max_frac=max_frac,
                                          sampler=bb.SobolQMCNormalSampler(torch.Size([256]), seed=9))  # fmt: skip
        idx = oracle_prune(ogps, X, cons, 0, 1.0, num_samples=256, seed=9, max_frac=max_frac)
        assert torch.equal(kept.cpu(), X[idx]) and 0 < kept.shape[0] < 60
    with pytest.raises(ValueError):
        prune_inferior_points_list(model, X.to(cuda), obj, max_frac=0.0)
    acqf = bb.qLogNoisyExpectedImprovement(model, X.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([32]), seed=1),
                                           objective=obj, constraints=cons, eta=1e-2, prune_baseline=True)  # fmt: skip
    assert acqf.X_baseline.shape[0] < 60
    pruned = acqf.X_baseline.cpu()
    oracle = OracleConstrainedQLogNEI(ogps, pruned, cons, 1e-2, obj_idx=0, obj_weight=1.0, num_samples=32, seed=1)
    Xc = torch.rand(3, 2, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(8))
    _check(cuda, acqf, oracle, Xc)


def test_constrained_qnei_matches_oracle(cuda):
    """qNoisyExpectedImprovement over a model list with constraints: relu improvement x sigmoid weights, amax, mean.  The
    baseline is a small subset of the training inputs, so that candidates do improve on the incumbent."""
    import warnings

    import botorch_b200 as bb

    X, ys, hy = _data(50, 3, seed=31)
    cons = [lambda Z: Z[..., 1] - 0.7]
    eta = torch.tensor([0.2], dtype=torch.float64)
    Xb = X[ys[0].argsort()[:8]]  # the eight worst observed points
    gms = [bb.SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["ls"], noise=h["noise"], mean_constant=h["mean"])
           for y, h in zip(ys[:2], hy[:2])]  # fmt: skip
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        acqf = bb.qNoisyExpectedImprovement(
            bb.ModelListGP(*gms), Xb.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([64]), seed=6),
            objective=bb.LinearMCObjective(torch.tensor([1.0, 0.0], dtype=torch.float64, device=cuda)), constraints=cons,
            eta=eta, prune_baseline=False)  # fmt: skip
    assert type(acqf).__name__ == "qNEIModelList" and not acqf._log
    ogps = [OracleGP(X, y.unsqueeze(-1), GPHyper("rbf", h["ls"], torch.tensor(h["noise"], dtype=torch.float64),
                                                 torch.tensor(h["mean"], dtype=torch.float64), None), standardize=True)
            for y, h in zip(ys[:2], hy[:2])]  # fmt: skip
    oracle = OracleConstrainedQLogNEI(ogps, Xb, cons, eta, obj_idx=0, obj_weight=1.0, num_samples=64, seed=6, fat=False,
                                      plain=True)  # fmt: skip
    Xc = torch.rand(6, 3, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(4))
    Xg = Xc.to(cuda).requires_grad_(True)
    val = acqf(Xg)
    (grad,) = torch.autograd.grad(val.sum(), Xg)
    oval, ograd = oracle.value_and_grad(Xc)
    assert bool((oval > 0).any())  # a non-trivial case: some candidates improve
    assert rel_err(val, oval, scale=float(oval.abs().max())) < TOL
    assert rel_err(grad, ograd, scale=float(ograd.abs().max())) < TOL


def test_constrained_path_reports_the_jitter_warning(cuda):
    """A candidate on a training point makes the q x q Schur block of every outcome singular: the jitter schedule runs in the
    kernels and check_numerics() reports it, as on the single-model path."""
    import warnings

    import botorch_b200 as bb

    X, ys, hy = _data(30, 3, seed=2)
    gms = [bb.SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["ls"], noise=1e-4, mean_constant=h["mean"])
           for y, h in zip(ys[:2], hy[:2])]  # fmt: skip
    acqf = bb.qLogNoisyExpectedImprovement(
        bb.ModelListGP(*gms), X.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([16]), seed=0),
        objective=bb.LinearMCObjective(torch.tensor([1.0, 0.0], dtype=torch.float64, device=cuda)),
        constraints=[lambda Z: Z[..., 1] - 0.9], eta=0.1)  # fmt: skip
    Xc = X[:2].clone().unsqueeze(0).to(cuda)  # two candidates sitting on training points
    Xc[0, 1] = Xc[0, 0]  # ... and on each other
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        val = acqf(Xc)
        acqf.check_numerics()
    assert bool(torch.isfinite(val).all()) or bool(torch.isneginf(val).all())
    assert any(issubclass(w.category, (bb.exceptions.NumericalWarning, bb.exceptions.BotorchWarning)) for w in caught)
