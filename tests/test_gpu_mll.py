"""GPU parity of the exact marginal log likelihood + hyperparameter gradient (bo_mll_value_grad) against
the oracle restatement (oracle/mll.py, torch autograd in float64), and of fit_gpytorch_mll against SciPy on the oracle."""
import math

import numpy as np
import pytest
import torch

from oracle import mll as omll
from tests.util import rel_err

pytestmark = pytest.mark.gpu

TOL = 1e-9  # north-star float64 tolerance


def _data(n, d, seed=0, noise=0.05):
    g = torch.Generator().manual_seed(seed)
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    y = torch.sin(6.0 * X[:, 0]) + 0.5 * X[:, -1] ** 2 + noise * torch.randn(n, generator=g, dtype=torch.float64)
    if n > 1:
        y = (y - y.mean()) / y.std()
    return X, y


def _hypers(d, seed, with_os):
    g = torch.Generator().manual_seed(100 + seed)
    h = {
        "lengthscale": 0.3 + 1.5 * torch.rand(d, generator=g, dtype=torch.float64),
        "noise": torch.tensor(0.03, dtype=torch.float64),
        "mean": torch.tensor(0.2, dtype=torch.float64),
        "outputscale": torch.tensor(1.7, dtype=torch.float64) if with_os else None,
    }
    return h


def _device_eval(cuda, X, y, kernel, h, noise_vec=None):
    from botorch_b200.engine import MLLState

    st = MLLState(X.to(cuda), y.to(cuda), kernel, with_outputscale=h["outputscale"] is not None,
                  noise_vec=None if noise_vec is None else noise_vec.to(cuda))  # fmt: skip
    v, g = st.evaluate(h["lengthscale"].to(cuda), h["noise"].to(cuda), h["mean"].to(cuda),
                       None if h["outputscale"] is None else h["outputscale"].to(cuda))  # fmt: skip
    torch.cuda.synchronize()
    return v.cpu(), g.cpu(), st


@pytest.mark.parametrize(
    "n,d,kernel,with_os",
    [(20, 2, "rbf", False), (50, 3, "rbf", True), (129, 5, "matern52", True), (300, 6, "rbf", False),
     (200, 20, "matern52", False), (1, 2, "rbf", False), (65, 1, "matern52", True), (700, 12, "rbf", True)],
)  # fmt: skip
def test_mll_value_and_gradient_match_oracle(cuda, n, d, kernel, with_os):
    X, y = _data(n, d, seed=n)
    h = _hypers(d, n, with_os)
    v, g, _ = _device_eval(cuda, X, y, kernel, h)
    ov, og = omll.value_and_grad(omll.log_marginal, X, y, kernel, h)
    assert abs(float(v[0]) - float(ov)) <= TOL * max(abs(float(ov)), 1.0)
    scale = max(float(og["lengthscale"].abs().max()), 1e-3 * n)  # the entries are sums of n^2 mixed-sign terms
    assert rel_err(g[0, :d], og["lengthscale"], scale=scale) < TOL
    assert abs(float(g[0, d + 1]) - float(og["noise"])) <= TOL * max(abs(float(og["noise"])), float(n))
    assert abs(float(g[0, d + 2]) - float(og["mean"])) <= TOL * max(abs(float(og["mean"])), 1.0)
    if with_os:
        assert abs(float(g[0, d]) - float(og["outputscale"])) <= TOL * max(abs(float(og["outputscale"])), float(n))


def test_mll_fixed_noise_vector(cuda):
    n, d = 90, 4
    X, y = _data(n, d, seed=5)
    h = _hypers(d, 5, True)
    nv = 0.01 + 0.05 * torch.rand(n, dtype=torch.float64, generator=torch.Generator().manual_seed(9))
    v, g, _ = _device_eval(cuda, X, y, "matern52", h, noise_vec=nv)
    ov, og = omll.value_and_grad(omll.log_marginal, X, y, "matern52", h, noise_vec=nv)
    assert abs(float(v[0]) - float(ov)) <= TOL * abs(float(ov))
    assert rel_err(g[0, :d], og["lengthscale"], scale=max(float(og["lengthscale"].abs().max()), 0.1)) < TOL
    assert abs(float(g[0, d]) - float(og["outputscale"])) <= TOL * n
    assert abs(float(g[0, d + 2]) - float(og["mean"])) <= TOL * max(abs(float(og["mean"])), 1.0)


def test_mll_leaves_valid_prediction_caches(cuda):
    """After an evaluation the GPState behind the MLL is factored for those hyperparameters."""
    from botorch_b200.engine import GPState

    n, d = 150, 3
    X, y = _data(n, d, seed=2)
    h = _hypers(d, 2, False)
    _, _, st = _device_eval(cuda, X, y, "rbf", h)
    ref = GPState(X.to(cuda), y.to(cuda), "rbf", h["lengthscale"].to(cuda), h["noise"].to(cuda), h["mean"].to(cuda))
    assert rel_err(st.gp.L, ref.L) < 1e-13
    assert rel_err(st.gp.alpha, ref.alpha) < 1e-12


def _model(cuda, X, y, **kw):
    from botorch_b200 import SingleTaskGP

    return SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), **kw)


def test_exact_mll_object_matches_oracle_with_priors(cuda):
    from botorch_b200.fit import ExactMarginalLogLikelihood, get_parameters_and_bounds

    n, d = 80, 4
    X, y = _data(n, d, seed=3)
    h = _hypers(d, 3, False)
    model = _model(cuda, X, 3.0 * y + 1.0, lengthscale=h["lengthscale"], noise=h["noise"], mean_constant=h["mean"])
    mll = ExactMarginalLogLikelihood(model.likelihood, model)
    names = [n_ for n_, _ in mll.named_parameters()]
    assert names == ["likelihood.noise_covar.raw_noise", "model.covar_module.raw_lengthscale", "model.mean_module.raw_constant"]
    params, bounds = get_parameters_and_bounds(mll)
    assert bounds["model.covar_module.raw_lengthscale"][0] == 2.5e-2 and bounds["likelihood.noise_covar.raw_noise"][0] == 1e-4
    assert "model.mean_module.raw_constant" not in bounds
    val, grads = mll.value_and_grads()
    targets = model.train_targets.cpu()  # standardised by the model's outcome transform
    ov, og = omll.value_and_grad(omll.exact_mll, X, targets, "rbf", h)
    assert abs(float(val) - float(ov)) <= TOL * abs(float(ov))
    assert rel_err(grads["lengthscale"].reshape(-1), og["lengthscale"], scale=float(og["lengthscale"].abs().max())) < TOL
    assert abs(float(grads["noise"]) - float(og["noise"])) <= TOL * max(abs(float(og["noise"])), 1.0)
    assert abs(float(grads["mean"]) - float(og["mean"])) <= TOL * max(abs(float(og["mean"])), 1.0)


def _oracle_fit(X, y, h0, maxiter=200):
    from scipy.optimize import minimize

    d = X.shape[-1]

    def f(x):
        h = {"lengthscale": torch.tensor(x[:d]), "noise": torch.tensor(x[d]), "mean": torch.tensor(x[d + 1]), "outputscale": None}
        v, g = omll.value_and_grad(omll.exact_mll, X, y, "rbf", h)
        return -float(v), -np.concatenate([g["lengthscale"].numpy(), [float(g["noise"])], [float(g["mean"])]])

    x0 = np.concatenate([h0["lengthscale"].numpy(), [float(h0["noise"])], [float(h0["mean"])]])
    bounds = [(2.5e-2, np.inf)] * d + [(1e-4, np.inf), (-np.inf, np.inf)]
    return minimize(f, x0, jac=True, bounds=bounds, method="L-BFGS-B", options={"maxiter": maxiter})


def test_fit_gpytorch_mll_reaches_the_oracle_optimum(cuda):
    from botorch_b200.fit import ExactMarginalLogLikelihood, fit_gpytorch_mll

    n, d = 60, 3
    X, y = _data(n, d, seed=7, noise=0.1)
    model = _model(cuda, X, y)
    h0 = {"lengthscale": model.covar_module.raw_lengthscale.detach().cpu().reshape(-1).clone(),
          "noise": model.likelihood.noise_covar.raw_noise.detach().cpu().reshape(()).clone(),
          "mean": model.mean_module.raw_constant.detach().cpu().clone()}  # fmt: skip
    mll = ExactMarginalLogLikelihood(model.likelihood, model)
    start = float(mll.value_and_grads()[0])
    out = fit_gpytorch_mll(mll)
    assert out is mll and not mll.training and not model.training
    fitted = float(mll.value_and_grads()[0])
    ref = _oracle_fit(X, model.train_targets.cpu(), h0)
    assert fitted > start
    # same objective, same optimiser, same start: identical gradients to 1e-9 give the same optimum up to L-BFGS-B's ftol
    assert abs(fitted - (-ref.fun)) < 1e-6 * max(abs(ref.fun), 1.0)
    ls = model.covar_module.raw_lengthscale.detach().cpu().reshape(-1)
    assert torch.allclose(ls, torch.tensor(ref.x[:d]), rtol=5e-3, atol=1e-4)
    # the fitted model serves posteriors on the fused path
    post = model.posterior(X[:5].to(cuda).unsqueeze(0))
    assert post.mean.shape == (1, 5, 1) and bool(torch.isfinite(post.variance).all())


def test_fit_that_hits_maxiter_is_returned_not_retried(cuda):
    """fit.py:51-70 (`_debug_warn`): L-BFGS-B stopping at its iteration limit is not a failed attempt -- the reference returns
    the partially fitted model; no ModelFittingError, no restart from prior samples."""
    import warnings

    from botorch_b200.fit import ExactMarginalLogLikelihood, fit_gpytorch_mll

    X, y = _data(40, 3, seed=5, noise=0.1)
    model = _model(cuda, X, y)
    mll = ExactMarginalLogLikelihood(model.likelihood, model)
    start = float(mll.value_and_grads()[0])
    evals0 = mll._state.evaluations if getattr(mll, "_state", None) is not None else None
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        out = fit_gpytorch_mll(mll, optimizer_kwargs={"options": {"maxiter": 1}})
    assert out is mll and not mll.training
    assert float(mll.value_and_grads()[0]) > start  # one accepted step
    if evals0 is not None:
        assert mll._state.evaluations - evals0 < 25  # a single attempt, not five


def test_fit_retries_and_raises_model_fitting_error(cuda):
    from botorch_b200.exceptions import OptimizationWarning
    from botorch_b200.fit import ExactMarginalLogLikelihood, ModelFittingError, OptimizationResult, OptimizationStatus, fit_gpytorch_mll

    X, y = _data(30, 2, seed=11)
    model = _model(cuda, X, y)
    mll = ExactMarginalLogLikelihood(model.likelihood, model)
    calls = []

    def failing(mll_, closure=None, **kw):
        import warnings

        calls.append(model.covar_module.raw_lengthscale.detach().cpu().clone())
        warnings.warn("did not converge", OptimizationWarning)
        return OptimizationResult(step=0, fval=float("nan"), status=OptimizationStatus.FAILURE)

    with pytest.raises(ModelFittingError), pytest.warns(OptimizationWarning):
        fit_gpytorch_mll(mll, optimizer=failing, max_attempts=3)
    assert len(calls) == 3 and not torch.equal(calls[0], calls[1])  # attempts 2.. start from prior samples
    assert mll.training
    assert torch.equal(model.covar_module.raw_lengthscale.detach().cpu(), calls[0])  # rolled back


def test_mll_batched_hyperparameter_sets(cuda):
    """M hyperparameter sets over one training set in a single call (the shape of a multi-chain NUTS potential for the
    SAAS model, models/fully_bayesian.py:168-191): every member matches its own oracle evaluation."""
    from botorch_b200.engine import MLLState

    n, d, M = 100, 30, 5
    X, y = _data(n, d, seed=21)
    g = torch.Generator().manual_seed(4)
    ls = 0.2 + 2.0 * torch.rand(M, d, generator=g, dtype=torch.float64)
    noise = 1e-3 + 0.1 * torch.rand(M, generator=g, dtype=torch.float64)
    mean = 0.3 * torch.randn(M, generator=g, dtype=torch.float64)
    os_ = 0.5 + 1.5 * torch.rand(M, generator=g, dtype=torch.float64)
    st = MLLState(X.to(cuda), y.to(cuda), "matern52", with_outputscale=True, M=M)
    v, gr = st.evaluate(ls.to(cuda), noise.to(cuda), mean.to(cuda), os_.to(cuda))
    v, gr = v.cpu(), gr.cpu()
    for m in range(M):
        h = {"lengthscale": ls[m], "noise": noise[m], "mean": mean[m], "outputscale": os_[m]}
        ov, og = omll.value_and_grad(omll.log_marginal, X, y, "matern52", h)
        assert abs(float(v[m]) - float(ov)) <= TOL * abs(float(ov))
        assert rel_err(gr[m, :d], og["lengthscale"], scale=max(float(og["lengthscale"].abs().max()), 0.1)) < TOL
        assert abs(float(gr[m, d]) - float(og["outputscale"])) <= TOL * n
        assert abs(float(gr[m, d + 1]) - float(og["noise"])) <= TOL * max(abs(float(og["noise"])), float(n))
        assert abs(float(gr[m, d + 2]) - float(og["mean"])) <= TOL * max(abs(float(og["mean"])), 1.0)


def test_sum_mll_fits_every_outcome_of_a_model_list(cuda):
    """fit_gpytorch_mll(SumMarginalLogLikelihood(model.likelihood, model)) == fitting each outcome on its own (_fit_list,
    fit.py:262-285), also for a SingleTaskGP built on n x m targets."""
    import botorch_b200 as bb

    n, d = 50, 3
    X, y0 = _data(n, d, seed=3, noise=0.1)
    y1 = torch.cos(4.0 * X[:, 1]) + 0.2 * torch.randn(n, generator=torch.Generator().manual_seed(4), dtype=torch.float64)
    Y = torch.stack([y0.reshape(-1), y1.reshape(-1)], dim=-1)
    multi = bb.SingleTaskGP(X.to(cuda), Y.to(cuda))
    assert isinstance(multi, bb.ModelListGP)
    mll = bb.SumMarginalLogLikelihood(multi.likelihood, multi)
    before = float(mll.forward())
    out = bb.fit_gpytorch_mll(mll)
    assert out is mll and not mll.training and float(mll.forward()) > before
    for k in range(2):
        single = bb.SingleTaskGP(X.to(cuda), Y[:, k : k + 1].to(cuda))
        bb.fit_gpytorch_mll(bb.ExactMarginalLogLikelihood(single.likelihood, single))
        a, b = multi.models[k].state_dict(), single.state_dict()
        assert a.keys() == b.keys() and all(torch.equal(a[key], b[key]) for key in a)
    post = multi.posterior(X[:4].to(cuda).unsqueeze(0))
    assert post.mean.shape == (1, 4, 2) and bool(torch.isfinite(post.variance).all())
