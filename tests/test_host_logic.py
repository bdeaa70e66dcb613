"""CPU tests of the host-side logic (no compute calls into the CUDA library): C-ABI symbols, samplers and their
base-sample bookkeeping, initial-condition selection, shard arithmetic, error behaviour without a GPU."""
import ctypes
import os
import re
import warnings

import pytest
import torch

import botorch_b200 as bb
from botorch_b200 import distributed as bdist
from botorch_b200 import exceptions as bex
from botorch_b200 import optim as boptim
from botorch_b200.acquisition import _ShapeOnlyPosterior
from botorch_b200.build import LIB_PATH, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_builds_and_exports_every_declared_symbol():
    path = build()
    assert path == LIB_PATH and os.path.exists(path)
    lib = ctypes.CDLL(path)
    header = open(os.path.join(ROOT, "include", "botorch_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    names = set(re.findall(r"\b(bo_[a-z0-9_]+)\s*\(", header))
    assert len(names) >= 25
    missing = [n for n in sorted(names) if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.bo_pad8(13) == 16 and lib.bo_kp(20) == 24
    lib.bo_strerror.restype = ctypes.c_char_p
    assert b"workspace" in lib.bo_strerror(3)


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    X, Y = torch.rand(6, 2, dtype=torch.float64), torch.rand(6, 1, dtype=torch.float64)
    model = bb.SingleTaskGP(X, Y)
    with pytest.raises(Exception) as ei:
        model.posterior(torch.rand(3, 1, 2, dtype=torch.float64))
    assert "CPU" in str(ei.value) or "CUDA" in str(ei.value)
    with pytest.raises(Exception):
        bb.qLogNoisyExpectedImprovement(model, X)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "botorch_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f


def test_model_argument_validation_and_state_dict():
    X, Y = torch.rand(5, 3, dtype=torch.float64), torch.rand(5, 1, dtype=torch.float64)
    with pytest.raises(bex.BotorchTensorDimensionError):
        bb.SingleTaskGP(X, Y.squeeze(-1))
    with pytest.raises(bex.InputDataError):
        bb.SingleTaskGP(X, Y.float())
    with pytest.warns(bex.InputDataWarning):
        bb.SingleTaskGP(X.float(), Y.float())
    with pytest.raises(bex.UnsupportedError):
        bb.SingleTaskGP(X, Y, covar_module=object())
    with pytest.raises(bex.BotorchTensorDimensionError):  # Yvar needs the explicit output dimension (gpytorch.py:91-127)
        bb.SingleTaskGP(X, Y, train_Yvar=torch.ones(5, dtype=torch.float64))
    with pytest.warns(bex.NumericalWarning):  # gpytorch's min_fixed_noise floor
        fixed = bb.SingleTaskGP(X, Y, train_Yvar=torch.full_like(Y, 1e-9))
    assert float(fixed.train_Yvar.min()) == 1e-6
    m = bb.SingleTaskGP(X, Y)
    sd = m.state_dict()
    for k in ("covar_module.raw_lengthscale", "likelihood.noise_covar.raw_noise", "mean_module.raw_constant",
              "outcome_transform.means", "outcome_transform.stdvs"):  # fmt: skip
        assert k in sd, k
    assert sd["covar_module.raw_lengthscale"].shape == (1, 3)
    assert torch.allclose(sd["outcome_transform.means"].reshape(()), Y.mean())
    assert torch.allclose(m.train_targets.std(), torch.tensor(1.0, dtype=torch.float64))
    sd2 = {k: v.clone() for k, v in sd.items()}
    sd2["covar_module.raw_lengthscale"] = torch.full((1, 3), 0.7, dtype=torch.float64)
    sd2["covar_module.raw_lengthscale_constraint.lower_bound"] = torch.tensor(0.025)  # extra reference key: ignored
    m.load_state_dict(sd2)
    assert torch.allclose(m._hypers()["lengthscale"], torch.full((3,), 0.7, dtype=torch.float64))
    saas = bb.SaasFullyBayesianSingleTaskGP(X, Y)
    with pytest.raises(RuntimeError):
        saas.posterior(X)
    saas.load_mcmc_samples({"mean": torch.zeros(4), "lengthscale": torch.ones(4, 3), "outputscale": torch.ones(4),
                            "noise": torch.full((4,), 1e-6)})  # fmt: skip
    assert saas.batch_shape == torch.Size([4]) and saas.num_mcmc_samples == 4
    assert float(saas._hypers()["noise"].min()) == pytest.approx(1e-4)  # MIN_INFERRED_NOISE_LEVEL clamp
    sd = saas.state_dict()
    assert sd["covar_module.base_kernel.raw_lengthscale"].shape == (4, 1, 3)
    saas2 = bb.SaasFullyBayesianSingleTaskGP(X, Y)
    saas2.load_state_dict(sd)
    assert torch.allclose(saas2._hypers()["lengthscale"], saas._hypers()["lengthscale"])
    saas.train()
    with pytest.raises(RuntimeError):
        saas.num_mcmc_samples


def test_sampler_shapes_seeding_and_baseline_columns():
    """test/sampling/test_normal.py:75-116 + sampling/normal.py:68-131 on shape-only posteriors."""
    S = torch.Size([8])
    with pytest.raises(ValueError):
        bb.SobolQMCNormalSampler(sample_shape=8)
    post_b = _ShapeOnlyPosterior(torch.Size(), 5, torch.device("cpu"), torch.float64)
    s1 = bb.SobolQMCNormalSampler(S, seed=1234)
    s1._construct_base_samples(post_b)
    assert s1.base_samples.shape == (8, 5)
    first = s1.base_samples.clone()
    s1._construct_base_samples(post_b)  # same shape: no redraw
    assert torch.equal(s1.base_samples, first)
    s2 = bb.SobolQMCNormalSampler(S, seed=1234)
    s2._construct_base_samples(post_b)
    assert torch.equal(s2.base_samples, first)  # seeded determinism
    s3 = bb.SobolQMCNormalSampler(S, seed=1235)
    s3._construct_base_samples(post_b)
    assert not torch.equal(s3.base_samples, first)
    # appending q = 3 points: t-batch collapsed, baseline columns preserved, same seed for the fresh draw
    from copy import deepcopy

    base = deepcopy(s1)
    post_f = _ShapeOnlyPosterior(torch.Size([1]), 8, torch.device("cpu"), torch.float64)
    s1._update_base_samples(post_f, base)
    assert s1.base_samples.shape == (8, 1, 8)
    assert torch.equal(s1.base_samples[:, 0, :5], first)
    fresh = bb.draw_sobol_normal_samples(d=8, n=8, dtype=torch.float64, seed=1234)
    assert torch.equal(s1.base_samples[:, 0, 5:], fresh[:, 5:])
    # ensemble model: model batch dim collapsed too (posteriors/fully_bayesian.py:150-158)
    e = bb.SobolQMCNormalSampler(S, seed=3)
    e._construct_base_samples(_ShapeOnlyPosterior(torch.Size([6]), 5, torch.device("cpu"), torch.float64))
    assert e.base_samples.shape == (8, 1, 5)
    eb = deepcopy(e)
    e._update_base_samples(_ShapeOnlyPosterior(torch.Size([1, 6]), 7, torch.device("cpu"), torch.float64), eb)
    assert e.base_samples.shape == (8, 1, 1, 7) and torch.equal(e.base_samples[:, 0, 0, :5], eb.base_samples[:, 0])
    with pytest.raises(RuntimeError):
        e._update_base_samples(post_f, bb.IIDNormalSampler(S, seed=0))
    iid = bb.IIDNormalSampler(S, seed=5)
    iid._construct_base_samples(post_b)
    iid2 = bb.IIDNormalSampler(S, seed=5)
    iid2._construct_base_samples(post_b)
    assert torch.equal(iid.base_samples, iid2.base_samples)
    big = _ShapeOnlyPosterior(torch.Size(), 30000, torch.device("cpu"), torch.float64)
    assert isinstance(bb.get_sampler(big, S), bb.IIDNormalSampler)
    assert isinstance(bb.get_sampler(post_b, S), bb.SobolQMCNormalSampler)
    with pytest.raises(bex.UnsupportedError):
        bb.SobolQMCNormalSampler(S)._construct_base_samples(big)


def test_sobol_draw_is_bit_identical_to_the_stock_engine():
    """``sampling.sobol_draw`` (column-sliced, split by dimension over host threads) against torch's ``SobolEngine`` --
    the generator the reference's samplers use (sampling/qmc.py:39-105, utils/sampling.py:97-140): ``torch.equal``."""
    from torch.quasirandom import SobolEngine

    from botorch_b200.sampling import NormalQMCEngine, sobol_draw

    for dim, n, dtype in [(1, 1, torch.float64), (7, 5, torch.float64), (160, 1024, torch.float64), (408, 64, torch.float32),
                          (1203, 33, torch.float64)]:  # fmt: skip
        ref = SobolEngine(dim, scramble=True, seed=99).draw(n, dtype=dtype)
        assert torch.equal(sobol_draw(dim, n, seed=99, dtype=dtype), ref)
        cols = torch.arange(dim)[-(dim // 3 + 1) :]
        assert torch.equal(sobol_draw(dim, n, seed=99, dtype=dtype, columns=cols), ref[:, cols])
        cols = torch.nonzero(torch.arange(dim) % 5 >= 3).squeeze(-1)  # strided selection (ensemble-collapsed layout)
        assert torch.equal(sobol_draw(dim, n, seed=99, dtype=dtype, columns=cols), ref[:, cols])
    # the box map and the inverse normal CDF are applied per column block by the workers: same roundings as the reference's
    # `lower + width * u` / `erfinv(2 v - 1) sqrt 2` over the whole tensor
    bounds = torch.tensor([[-1.0, 0.5, 2.0], [1.0, 0.75, 5.0]], dtype=torch.float64)
    u = SobolEngine(12, scramble=True, seed=5).draw(1000, dtype=torch.float64).view(1000, 4, 3)
    ref = bounds[0] + (bounds[1] - bounds[0]) * u
    assert torch.equal(bb.draw_sobol_samples(bounds, 1000, 4, seed=5), ref)
    assert torch.equal(bb.draw_sobol_samples(bounds, 1000, 4, seed=5, rows=(100, 300)), ref[100:300])
    u = SobolEngine(300, scramble=True, seed=2).draw(64, dtype=torch.float64)
    v = 0.5 + (1 - torch.finfo(u.dtype).eps) * (u - 0.5)
    assert torch.equal(bb.draw_sobol_normal_samples(300, 64, dtype=torch.float64, seed=2), torch.erfinv(2 * v - 1) * 2**0.5)
    # unseeded: the scramble bits come from the global stream in the same order
    torch.manual_seed(7)
    ref = SobolEngine(40, scramble=True).draw(16, dtype=torch.float64)
    torch.manual_seed(7)
    assert torch.equal(sobol_draw(40, 16, dtype=torch.float64), ref)
    with pytest.raises(ValueError):
        sobol_draw(SobolEngine.MAXDIM + 1, 2, seed=0)
    # the normal engine keeps its position across draws
    e = NormalQMCEngine(d=6, seed=5, inv_transform=True)
    a, b = e.draw(4, dtype=torch.float64), e.draw(3, dtype=torch.float64)
    u = SobolEngine(6, scramble=True, seed=5).draw(7, dtype=torch.float64)
    v = 0.5 + (1 - torch.finfo(u.dtype).eps) * (u - 0.5)
    assert torch.equal(torch.cat([a, b]), torch.erfinv(2 * v - 1) * 2**0.5)
    # appended points of the cached-Cholesky path: only the last q Sobol dimensions are computed
    from copy import deepcopy

    S, n_b, q = torch.Size([16]), 300, 4
    s = bb.SobolQMCNormalSampler(S, seed=11)
    s._construct_base_samples(_ShapeOnlyPosterior(torch.Size(), n_b, torch.device("cpu"), torch.float64))
    base = deepcopy(s)
    s._update_base_samples(_ShapeOnlyPosterior(torch.Size([1]), n_b + q, torch.device("cpu"), torch.float64), base)
    u = SobolEngine(n_b + q, scramble=True, seed=11).draw(16, dtype=torch.float64)
    v = 0.5 + (1 - torch.finfo(u.dtype).eps) * (u - 0.5)
    assert torch.equal(s.base_samples[:, 0, n_b:], (torch.erfinv(2 * v - 1) * 2**0.5)[:, n_b:])
    assert torch.equal(s.base_samples[:, 0, :n_b], base.base_samples)


def test_initialize_q_batch_semantics():
    """optim/initializers.py:893-965 and test/optim/test_initializers.py:131-243."""
    torch.manual_seed(0)
    X = torch.rand(40, 2, 3, dtype=torch.float64)
    Y = torch.randn(40, dtype=torch.float64)
    out = bb.initialize_q_batch(X, Y, n=5)
    assert out.shape == (5, 2, 3)
    best = X[int(torch.argmax(Y))]
    assert any(torch.equal(o, best) for o in out)  # the argmax is always kept
    assert bb.initialize_q_batch(X, Y, n=40) is X
    with pytest.raises(RuntimeError):
        bb.initialize_q_batch(X, Y, n=41)
    with pytest.warns(bex.BadInitialCandidatesWarning):
        r = bb.initialize_q_batch(X, torch.zeros(40, dtype=torch.float64), n=5)
    assert r.shape == (5, 2, 3)
    huge = torch.linspace(0, 1, 40, dtype=torch.float64)
    huge[7] = 1e6  # exp overflow guard (eta halving)
    out = bb.initialize_q_batch(X, huge, n=3, eta=1e3)
    assert any(torch.equal(o, X[7]) for o in out)


def test_optimize_acqf_plumbing_with_a_host_function():
    """Control flow of optimize_acqf / gen_candidates_scipy on a closed-form acquisition (no GP)."""
    target = torch.tensor([0.3, 0.7], dtype=torch.float64)

    class Quad(torch.nn.Module):
        calls = 0

        def forward(self, X):
            Quad.calls += 1
            return -((X - target) ** 2).sum(dim=(-1, -2))

    bounds = torch.stack([torch.zeros(2), torch.ones(2)]).double()
    cand, val = bb.optimize_acqf(Quad(), bounds, q=1, num_restarts=6, raw_samples=32,
                                 options={"seed": 1, "batch_limit": 4, "init_batch_limit": 10, "maxiter": 50})  # fmt: skip
    assert cand.shape == (1, 2) and torch.allclose(cand[0], target, atol=1e-5) and float(val) > -1e-9
    c_all, v_all = bb.optimize_acqf(Quad(), bounds, q=2, num_restarts=3, raw_samples=8, options={"seed": 1},
                                    return_best_only=False)  # fmt: skip
    assert c_all.shape == (3, 2, 2) and v_all.shape == (3,)
    with pytest.raises(ValueError):
        bb.optimize_acqf(Quad(), bounds, q=1, num_restarts=2)
    with pytest.raises(bex.UnsupportedError):
        bb.optimize_acqf(Quad(), bounds, q=1, num_restarts=2, raw_samples=4, fixed_features={0: 0.5})
    # NaN from the acquisition function -> RuntimeError (gen.py:196-220)
    nan_f = lambda X: (X * float("nan")).sum(dim=(-1, -2))  # noqa: E731
    with pytest.raises(RuntimeError):
        bb.gen_candidates_scipy(torch.rand(2, 1, 2, dtype=torch.float64), nan_f, bounds[0], bounds[1])
    ics = torch.tensor([[[1.5, -0.2]]], dtype=torch.float64)  # out of bounds start is clamped
    c, v = bb.gen_candidates_scipy(ics, Quad(), bounds[0], bounds[1], options={"maxiter": 5})
    assert bool(((c >= 0) & (c <= 1)).all())
    assert torch.equal(bb.get_best_candidates(c_all, v_all), c_all[int(torch.argmax(v_all))])
    assert boptim.columnwise_clamp(ics, 0.0, 1.0).tolist() == [[[1.0, 0.0]]]
    with pytest.raises(RuntimeError):
        boptim.columnwise_clamp(ics, 0.0, 1.0, raise_on_violation=True)


def test_shard_range_partitions_exactly():
    for total in (0, 1, 7, 512, 32768, 1000003):
        for ws in (1, 2, 3, 4, 8):
            blocks = [bdist.shard_range(total, r, ws) for r in range(ws)]
            assert blocks[0][0] == 0 and blocks[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1
    assert bdist.world() == (0, 1)
    v = torch.tensor([1.0, 5.0, 5.0, 2.0], dtype=torch.float64)
    c = torch.arange(4, dtype=torch.float64).reshape(4, 1, 1)
    cand, val, idx = bdist.gather_best(v, c, index_offset=10)
    assert idx == 11 and float(val) == 5.0 and float(cand) == 1.0  # first occurrence wins


def test_expma_stopping_criterion():
    """optim/stopping.py:83-127 / test/optim/test_stopping.py: stops at maxiter, not before the window is full, and
    once the moving average of a flat objective stops decreasing."""
    from botorch_b200.optim import ExpMAStoppingCriterion

    sc = ExpMAStoppingCriterion(maxiter=2)
    assert not sc.evaluate(torch.tensor(1.0)) and sc.evaluate(torch.tensor(1.0))
    sc = ExpMAStoppingCriterion(n_window=4, rel_tol=0.0375)
    stops = [sc.evaluate(torch.tensor(1.0 - 0.01 * i, dtype=torch.float64)) for i in range(6)]
    assert stops[:4] == [False] * 4 and stops[4]  # relative decrease 0.01 per step < rel_tol
    sc = ExpMAStoppingCriterion(n_window=3, rel_tol=1e-6)
    vals = torch.tensor([[10.0, 5.0], [5.0, 4.0], [2.5, 3.9], [1.2, 3.9], [0.6, 3.9]], dtype=torch.float64)
    assert not any(sc.evaluate(v) for v in vals)  # one element still improving => keep going


def test_gen_candidates_torch_on_a_quadratic_and_host_evaluator_cpu():
    """The torch-optimiser loop and the SciPy host bridge are generic over the acquisition callable: maximise
    -(x - 0.3)^2 inside [0, 1] from several starts."""
    from botorch_b200.optim import HostEvaluator, gen_candidates_torch

    acq = lambda X: -((X - 0.3) ** 2).sum(dim=(-1, -2))
    ics = torch.rand(5, 2, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(0))
    cands, vals = gen_candidates_torch(ics, acq, lower_bounds=0.0, upper_bounds=1.0, options={"maxiter": 400, "lr": 0.05})
    assert cands.shape == ics.shape and vals.shape == (5,)
    assert float((cands - 0.3).abs().max()) < 2e-2 and float(vals.min()) > -1e-2
    ev = HostEvaluator(acq, ics.shape, ics)
    loss, grad = ev(ics.reshape(-1).numpy())
    assert abs(loss - float(-acq(ics).sum())) < 1e-12
    assert torch.allclose(torch.from_numpy(grad).view_as(ics), 2 * (ics - 0.3))


def test_thompson_winner_reduction_and_block_keyed_noise():
    """distributed._winner: across ranks the largest value wins, lowest global index on ties (torch.argmax semantics,
    generation/sampling.py:117); base samples of a candidate depend on its global block only."""
    from botorch_b200 import distributed as bd

    vals = torch.tensor([[1.0, 5.0, 2.0], [3.0, 5.0, -float("inf")]], dtype=torch.float64)  # ranks x draws
    idxs = torch.tensor([[10, 40, 7], [99, 12, torch.iinfo(torch.int64).max]])
    assert bd._winner(vals, idxs).tolist() == [99, 12, 7]
    z_all = bd.ts_base_samples(0, 40000, 1, 3, torch.device("cpu"))
    z_mid = bd.ts_base_samples(16000, 33000, 1, 3, torch.device("cpu"))
    assert torch.equal(z_all[16000:33000], z_mid)
    assert not torch.equal(z_all[:100], bd.ts_base_samples(0, 100, 2, 3, torch.device("cpu")))


# ---- fit.py host logic (no compute calls into the CUDA library) -----------------------------------------------------------
def _cpu_model(n=12, d=3):
    from botorch_b200 import SingleTaskGP

    g = torch.Generator().manual_seed(0)
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    return SingleTaskGP(X, torch.randn(n, 1, generator=g, dtype=torch.float64))


def test_fit_parameter_names_bounds_and_prior_resampling():
    from botorch_b200.fit import ExactMarginalLogLikelihood, get_parameters_and_bounds, sample_all_priors

    model = _cpu_model()
    mll = ExactMarginalLogLikelihood(model.likelihood, model)
    assert [k for k, _ in mll.named_parameters()] == [
        "likelihood.noise_covar.raw_noise", "model.covar_module.raw_lengthscale", "model.mean_module.raw_constant"]  # fmt: skip
    params, bounds = get_parameters_and_bounds(mll)
    assert set(params) == {"likelihood.noise_covar.raw_noise", "model.covar_module.raw_lengthscale", "model.mean_module.raw_constant"}
    assert bounds["model.covar_module.raw_lengthscale"] == (2.5e-2, float("inf"))
    assert bounds["likelihood.noise_covar.raw_noise"] == (1e-4, float("inf"))
    assert "model.mean_module.raw_constant" not in bounds  # ConstantMean carries no constraint
    before = model.covar_module.raw_lengthscale.detach().clone()
    torch.manual_seed(1)
    sample_all_priors(model)
    after = model.covar_module.raw_lengthscale.detach()
    assert not torch.equal(before, after) and bool((after >= 2.5e-2).all())
    assert float(model.likelihood.noise_covar.raw_noise) >= 1e-4


def test_scipy_minimize_drives_a_closure_over_named_tensors():
    from botorch_b200.fit import OptimizationStatus, scipy_minimize

    params = {"a": torch.nn.Parameter(torch.tensor([3.0, -2.0], dtype=torch.float64)),
              "b": torch.nn.Parameter(torch.tensor(5.0, dtype=torch.float64))}  # fmt: skip
    target = {"a": torch.tensor([1.0, 0.5], dtype=torch.float64), "b": torch.tensor(-1.0, dtype=torch.float64)}

    def closure():
        loss = sum(((p - target[k]) ** 2).sum() for k, p in params.items())
        grads = torch.autograd.grad(loss, list(params.values()))
        return loss.detach(), grads

    seen = []
    res = scipy_minimize(closure, params, bounds={"b": (0.0, None)}, callback=lambda p, r: seen.append(r.step))
    assert res.status == OptimizationStatus.SUCCESS and res.step >= 1 and seen == list(range(1, len(seen) + 1))
    assert torch.allclose(params["a"].detach(), target["a"], atol=1e-6)
    assert abs(float(params["b"])) < 1e-8  # the bound is active: optimum at the lower bound 0, not at -1
    with torch.no_grad():
        params["a"].copy_(torch.tensor([30.0, -20.0], dtype=torch.float64))
    stopped = scipy_minimize(closure, params, options={"maxiter": 1})  # iteration limit: STOPPED, not FAILURE
    assert stopped.status == OptimizationStatus.STOPPED and "ITERATIONS REACHED LIMIT" in stopped.message.upper()


def test_fit_gpytorch_mll_retry_policy_on_the_host():
    import warnings

    from botorch_b200.exceptions import OptimizationWarning
    from botorch_b200.fit import (ExactMarginalLogLikelihood, ModelFittingError, OptimizationResult, OptimizationStatus,
                                  fit_gpytorch_mll)  # fmt: skip

    model = _cpu_model()
    mll = ExactMarginalLogLikelihood(model.likelihood, model)
    start = model.covar_module.raw_lengthscale.detach().clone()
    calls = []

    def flaky(mll_, closure=None, **kw):  # warns on the first two attempts, succeeds on the third
        calls.append(model.covar_module.raw_lengthscale.detach().clone())
        if len(calls) < 3:
            warnings.warn("line search failed", OptimizationWarning)
        return OptimizationResult(step=1, fval=1.0, status=OptimizationStatus.SUCCESS)

    with pytest.warns(OptimizationWarning):
        out = fit_gpytorch_mll(mll, optimizer=flaky, closure=lambda: None, max_attempts=5)
    assert out is mll and not mll.training and len(calls) == 3
    assert torch.equal(calls[0], start) and not torch.equal(calls[1], start)  # attempts 2.. restart from prior samples

    def failing(mll_, closure=None, **kw):
        warnings.warn("did not converge", OptimizationWarning)
        return OptimizationResult(step=0, fval=float("nan"), status=OptimizationStatus.FAILURE)

    now = model.covar_module.raw_lengthscale.detach().clone()
    with pytest.raises(ModelFittingError), pytest.warns(OptimizationWarning):
        fit_gpytorch_mll(mll, optimizer=failing, closure=lambda: None, max_attempts=2)
    assert mll.training and torch.equal(model.covar_module.raw_lengthscale.detach(), now)  # rolled back


def test_fit_warning_handler_follows_the_reference_policy():
    """fit.py:51-70 + optim/utils/common.py:35-60: an L-BFGS-B iteration-limit message is resolved quietly (no retry, no
    ModelFittingError), a non-optimisation warning and the time-out message are re-emitted without a retry."""
    import warnings

    from botorch_b200.exceptions import OptimizationWarning
    from botorch_b200.fit import ExactMarginalLogLikelihood, OptimizationResult, OptimizationStatus, fit_gpytorch_mll

    model = _cpu_model()
    mll = ExactMarginalLogLikelihood(model.likelihood, model)
    calls = []

    def hits_maxiter(mll_, closure=None, **kw):
        calls.append(1)
        warnings.warn("`scipy_minimize` terminated with status OptimizationStatus.STOPPED, displaying original message "
                      "from `scipy.optimize.minimize`: STOP: TOTAL NO. of ITERATIONS REACHED LIMIT", OptimizationWarning)  # fmt: skip
        return OptimizationResult(step=1, fval=1.0, status=OptimizationStatus.STOPPED)

    with warnings.catch_warnings():
        warnings.simplefilter("error")  # nothing may be re-emitted
        out = fit_gpytorch_mll(mll, optimizer=hits_maxiter, closure=lambda: None, max_attempts=5)
    assert out is mll and not mll.training and len(calls) == 1
    calls.clear()

    def newer_scipy_wording(mll_, closure=None, **kw):
        calls.append(1)
        warnings.warn("STOP: TOTAL NO. OF F,G EVALUATIONS EXCEEDS LIMIT".replace("F,G", "f AND g"), OptimizationWarning)
        return OptimizationResult(step=1, fval=1.0, status=OptimizationStatus.STOPPED)

    with warnings.catch_warnings():
        warnings.simplefilter("error")
        fit_gpytorch_mll(mll, optimizer=newer_scipy_wording, closure=lambda: None, max_attempts=5)
    assert len(calls) == 1
    calls.clear()

    class NumericalWarning(RuntimeWarning):
        pass

    def jittered_and_timed_out(mll_, closure=None, **kw):
        calls.append(1)
        warnings.warn("A not p.d., added jitter of 1.0e-08 to the diagonal", NumericalWarning)
        warnings.warn("Optimization timed out after 1.0 seconds.", OptimizationWarning)
        return OptimizationResult(step=1, fval=1.0, status=OptimizationStatus.STOPPED)

    with pytest.warns() as rec:
        out = fit_gpytorch_mll(mll, optimizer=jittered_and_timed_out, closure=lambda: None, max_attempts=5)
    assert out is mll and len(calls) == 1  # both resolved: no retry
    cats = {r.category for r in rec}
    assert NumericalWarning in cats and OptimizationWarning in cats


# ---- NUTS host logic (botorch_b200/fully_bayesian.py): the sampler is device-agnostic, the potential is not ----------------
def test_nuts_recovers_a_gaussian_target():
    from botorch_b200.fully_bayesian import nuts_sample

    mu = torch.tensor([0.0, 1.0, -2.0], dtype=torch.float64)
    sd = torch.tensor([0.2, 1.0, 4.0], dtype=torch.float64)
    calls = []

    def potential(z):
        calls.append(z.shape)
        return 0.5 * (((z - mu) / sd) ** 2).sum(-1), (z - mu) / sd**2

    samples, stats = nuts_sample(potential, torch.zeros(6, 3, dtype=torch.float64), num_samples=250, warmup_steps=200, seed=3)
    flat = samples.reshape(-1, 3)
    assert samples.shape == (250, 6, 3) and all(s == (6, 3) for s in calls)
    assert bool(((flat.mean(0) - mu).abs() < 0.15 * sd).all())
    assert bool(((flat.std(0) / sd - 1.0).abs() < 0.12).all())
    assert 0.6 < float(stats["accept_prob"].mean()) < 0.99 and int(stats["divergent"].sum()) == 0
    # the adapted metric follows the target's scales (per chain, regularised variance of the warm-up windows)
    assert bool(((stats["inverse_mass"].sqrt() / sd).log().abs() < 0.7).all())
    again, _ = nuts_sample(potential, torch.zeros(6, 3, dtype=torch.float64), num_samples=5, warmup_steps=20, seed=3)
    again2, _ = nuts_sample(potential, torch.zeros(6, 3, dtype=torch.float64), num_samples=5, warmup_steps=20, seed=3)
    assert torch.equal(again, again2)


def test_nuts_treats_infinite_energy_as_divergence_and_keeps_the_chain():
    from botorch_b200.fully_bayesian import nuts_sample

    def potential(z):  # standard normal restricted to z > -1: a wall of infinite energy
        e = 0.5 * (z * z).sum(-1)
        bad = (z < -1.0).any(-1)
        return torch.where(bad, torch.full_like(e, float("inf")), e), torch.where(bad.unsqueeze(-1), torch.zeros_like(z), z)

    samples, stats = nuts_sample(potential, torch.zeros(4, 2, dtype=torch.float64), num_samples=150, warmup_steps=100, seed=0)
    assert bool((samples >= -1.0).all()) and bool(torch.isfinite(samples).all())
    assert int(stats["divergent"].sum()) > 0
    with pytest.raises(RuntimeError):
        nuts_sample(potential, torch.full((4, 2), -3.0, dtype=torch.float64), num_samples=2, warmup_steps=2)


def test_saas_prior_terms_and_adaptation_schedule():
    from torch.distributions import Gamma, HalfCauchy

    from botorch_b200 import fully_bayesian as fb

    x = torch.tensor([1e-3, 0.2, 1.0, 7.5], dtype=torch.float64)
    t = lambda v: torch.tensor(v, dtype=torch.float64)
    assert torch.allclose(fb._gamma_lp(x, x.log(), 2.0, 0.15), Gamma(t(2.0), t(0.15)).log_prob(x), rtol=1e-13, atol=0)
    assert torch.allclose(fb._gamma_lp(x, x.log(), 0.9, 10.0), Gamma(t(0.9), t(10.0)).log_prob(x), rtol=1e-13, atol=0)
    assert torch.allclose(fb._half_cauchy_lp(x, 0.1), HalfCauchy(t(0.1)).log_prob(x), rtol=1e-13, atol=0)
    start, ends = fb._adaptation_windows(512)
    assert start == 76 and ends[-1] == 460 and ends == sorted(ends) and ends[0] - start == 25
    assert fb._adaptation_windows(20) == (20, [])
    with pytest.raises(RuntimeError):
        fb.SaasPotential(torch.rand(5, 2, dtype=torch.float64), torch.rand(5, dtype=torch.float64))


def test_generic_objective_and_sample_path_dispatch_rules():
    """acquisition/objective.py:344-384: the callable receives (samples, X=X); which constructor arguments send an
    acquisition function to the sample-path classes of constrained.py (no device needed for the rule itself)."""
    from botorch_b200.acquisition import _needs_sample_path
    from botorch_b200.constrained import _objective_values, _one_outcome

    seen = {}

    def f(Y, X=None):
        seen["X"] = X
        return Y[..., 0] * Y[..., 1]

    obj = bb.GenericMCObjective(f)
    Y, X = torch.rand(5, 3, 2, 2, dtype=torch.float64), torch.rand(3, 2, 4, dtype=torch.float64)
    assert torch.equal(obj(Y, X=X), Y[..., 0] * Y[..., 1]) and seen["X"] is X
    assert _one_outcome(obj, 2) == (-1, 1.0)
    assert _one_outcome(None, 1) == (0, 1.0) and _one_outcome(bb.IdentityMCObjective(), 1) == (0, 1.0)
    assert _one_outcome(bb.LinearMCObjective(torch.tensor([0.0, -2.0])), 2) == (1, -2.0)
    with pytest.raises(bex.UnsupportedError):
        _one_outcome(None, 2)
    with pytest.raises(bex.UnsupportedError):
        _one_outcome(bb.LinearMCObjective(torch.tensor([1.0, 1.0])), 2)
    assert torch.equal(_objective_values(obj, -1, 1.0, Y, X), Y[..., 0] * Y[..., 1])
    assert torch.equal(_objective_values(None, 1, -2.0, Y, X), -2.0 * Y[..., 1])
    assert not _needs_sample_path(object(), None, None)
    assert not _needs_sample_path(object(), bb.LinearMCObjective(torch.tensor([1.0])), None)
    assert _needs_sample_path(object(), obj, None) and _needs_sample_path(object(), None, [lambda Z: Z[..., 0]])
