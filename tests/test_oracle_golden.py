"""CPU tests: the oracle against the golden vectors recorded from the reference's own modules
(tests/golden/reference_modules.pt, made by tests/golden/make_golden.py), its internal cross-checks, and the
known-answer / property tests of the reference's suite (test/utils/test_safe_math.py, test/acquisition/test_logei.py)
restated on the oracle."""
import math
import os

import pytest
import torch

from botorch_b200 import sampling as bsampling
from botorch_b200 import synthetic
from botorch_b200.synthetic import make_problem
from oracle import qmc, refstub, safe_math
from oracle.qlognei import OracleQLogNEI
from tests.util import candidates, oracle_acq, oracle_gp, rel_err

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def ref_gold():
    return torch.load(os.path.join(GOLD, "reference_modules.pt"))


@pytest.fixture(scope="module")
def cfg_gold():
    return torch.load(os.path.join(GOLD, "oracle_configs.pt"))


def _close(a, b, tol=1e-13):
    a, b = torch.as_tensor(a), torch.as_tensor(b)
    fin = torch.isfinite(b)
    assert torch.equal(torch.isfinite(a), fin)
    assert torch.equal(a[~fin], b[~fin]) or bool((torch.isnan(a[~fin]) == torch.isnan(b[~fin])).all())
    if fin.any():
        assert float(((a[fin] - b[fin]).abs() / b[fin].abs().clamp_min(1.0)).max()) <= tol


# ---- reductions pinned to botorch/utils/safe_math.py ------------------------------------------------------------------------
@pytest.mark.parametrize("tau", [1e-6, 1e-2, 1.0])
def test_log_fatplus_matches_reference(ref_gold, tau):
    c = ref_gold["safe_math"][f"log_fatplus_tau{tau:g}"]
    x = c["x"].clone().requires_grad_(True)
    v = safe_math.log_fatplus(x, tau=tau)
    (g,) = torch.autograd.grad(v.sum(), x)
    _close(v.detach(), c["value"])
    _close(g, c["grad"], 1e-12)
    _close(safe_math.fatplus(c["x"], tau=tau), ref_gold["safe_math"][f"fatplus_tau{tau:g}"]["value"])


@pytest.mark.parametrize("tau", [1e-2, 1.0])
def test_fatmax_matches_reference(ref_gold, tau):
    c = ref_gold["safe_math"][f"fatmax_tau{tau:g}"]
    x = c["x"].clone().requires_grad_(True)
    v = safe_math.fatmax(x, dim=-1, tau=tau)
    (g,) = torch.autograd.grad(v.sum(), x)
    _close(v.detach(), c["value"])
    _close(g, c["grad"], 1e-12)


def test_logmeanexp_and_infinities_match_reference(ref_gold):
    sm = ref_gold["safe_math"]
    c = sm["logmeanexp_dim0"]
    x = c["x"].clone().requires_grad_(True)
    v = safe_math.logmeanexp(x, dim=0)
    (g,) = torch.autograd.grad(v.sum(), x)
    _close(v.detach(), c["value"])
    _close(g, c["grad"], 1e-12)
    for name, fn in (("fatmax_inf", lambda t: safe_math.fatmax(t, dim=-1, tau=1e-2)),
                     ("logsumexp_inf", lambda t: safe_math.logsumexp(t, dim=-1)),
                     ("logmeanexp_inf", lambda t: safe_math.logmeanexp(t, dim=-1))):  # fmt: skip
        c = sm[name]
        z = c["x"].clone().requires_grad_(True)
        v = fn(z)
        _close(v.detach(), c["value"])
        (g,) = torch.autograd.grad(v[torch.isfinite(v)].sum(), z, allow_unused=True)
        _close(g, c["grad"], 1e-12)


def test_qlognei_reduction_chain_matches_reference(ref_gold):
    c = ref_gold["safe_math"]["qlognei_chain"]
    obj = c["obj"].clone().requires_grad_(True)
    acq = safe_math.qlognei_reduction(obj, c["best_f"], 1e-6, 1e-2)
    (g,) = torch.autograd.grad(acq.sum(), obj)
    _close(acq.detach(), c["value"])
    _close(g, c["grad"], 1e-12)


def test_known_answers_of_test_logei(ref_gold):
    """test_logei.py:324-395 / :429-501: baseline sample 0, new sample 1 -> exp(qLogNEI) == 1; batch -> [1, 0]."""
    ka = ref_gold["known_answers"]
    s = ka["single"]
    best = s["baseline"].max(dim=-1).values.unsqueeze(-1)  # S x 1
    acq = safe_math.qlognei_reduction(s["new"], best, 1e-6, 1e-2)
    assert abs(float(acq.exp()) - s["exp_acq"]) < 1e-5
    b = ka["batch"]
    acq = safe_math.qlognei_reduction(b["new"], best, 1e-6, 1e-2)
    assert torch.allclose(acq.exp(), torch.tensor(b["exp_acq"], dtype=torch.float64), atol=1e-6)


# ---- closed-form properties of test/utils/test_safe_math.py --------------------------------------------------------------------
def test_safe_math_properties():
    torch.manual_seed(0)
    x = torch.rand(50, 9, dtype=torch.float64) + 0.1
    assert torch.allclose(safe_math.logmeanexp(x.log(), dim=0), x.mean(0).log(), atol=1e-14)  # :244-269
    z = torch.zeros(1, dtype=torch.float64, requires_grad=True)  # :291-300 _pareto(0)=1, p'(0) = -1 convention
    p = safe_math.pareto(z, alpha=2.0)
    assert float(p) == 1.0
    y = torch.randn(200, 6, dtype=torch.float64, requires_grad=True)
    for tau in (1e-3, 1e-2, 1.0):
        fm = safe_math.fatmax(y, dim=-1, tau=tau)
        assert bool(((fm - y.amax(-1)).abs() <= math.log(6) * tau + 1e-12).all())  # :360-364
        (g,) = torch.autograd.grad(fm.sum(), y)
        assert torch.allclose(g.sum(-1), torch.ones(200, dtype=torch.float64), atol=1e-12)  # :390-391
        assert bool((g >= 0).all())
    t = torch.linspace(-5, 5, 101, dtype=torch.float64)
    for tau in (1e-3, 1e-1):
        assert bool(((safe_math.fatplus(t, tau) - t.clamp_min(0)).abs() <= tau).all())  # :308-342
        assert bool((safe_math.fatplus(t, tau) > 0).all())


# ---- Sobol draws pinned to botorch/sampling/qmc.py + botorch/utils/sampling.py ------------------------------------------------
def test_sobol_draws_match_reference(ref_gold):
    q = ref_gold["qmc"]
    for d, n, seed in ((3, 8, 0), (21, 16, 1234), (204, 32, 1234), (108, 4, 7)):
        want = q[f"normal_d{d}_n{n}_seed{seed}"]
        assert torch.equal(qmc.sobol_normal(d, n, seed), want)  # oracle
        assert torch.equal(bsampling.draw_sobol_normal_samples(d, n, dtype=torch.float64, seed=seed), want)  # product
    bounds = torch.tensor([[0.0, -1.0, 2.0], [1.0, 1.0, 5.0]], dtype=torch.float64)
    assert torch.equal(qmc.sobol_box(bounds, 8, 2, 0), q["box_n8_q2_seed0"])
    assert torch.equal(bsampling.draw_sobol_samples(bounds, 8, 2, seed=0), q["box_n8_q2_seed0"])
    got = bsampling.draw_sobol_samples(bounds, 4, 1, batch_shape=torch.Size([3]), seed=5)
    assert torch.equal(got, q["box_n4_q1_batch3_seed5"])
    eng = bsampling.NormalQMCEngine(d=5, seed=11, inv_transform=False)
    assert torch.equal(eng.draw(6, dtype=torch.float64), q["boxmuller_d5_n6_seed11"])


def test_synthetic_functions_match_reference(ref_gold):
    s = ref_gold["synthetic"]
    _close(synthetic.branin(s["branin"]["X"]), s["branin"]["f"], 1e-14)
    _close(synthetic.hartmann6(s["hartmann6"]["X"]), s["hartmann6"]["f"], 1e-14)
    _close(synthetic.ackley(s["ackley20"]["X"]), s["ackley20"]["f"], 1e-14)


@pytest.mark.skipif(not refstub.available(), reason="/root/reference is not present on this machine")
def test_golden_file_is_reproducible_from_the_live_reference(ref_gold):
    ref = refstub.load()
    want = ref_gold["qmc"]["normal_d21_n16_seed1234"]
    assert torch.equal(ref.sampling.draw_sobol_normal_samples(d=21, n=16, dtype=torch.float64, seed=1234), want)
    x = ref_gold["safe_math"]["fatmax_tau0.01"]["x"]
    assert torch.equal(ref.safe_math.fatmax(x, dim=-1, tau=1e-2), ref_gold["safe_math"]["fatmax_tau0.01"]["value"])


# ---- the GP part of the oracle: two independent formulations + the reference's self-consistency identities -----------------
@pytest.mark.parametrize("cfg,kw,q", [("C1", {}, 2), ("C2", {"n": 80}, 4), ("C4", {"n": 40, "M": 3}, 3), ("C2", {"n": 50, "noise": 1e-4}, 4)])
def test_posterior_formulations_agree(cfg, kw, q):
    p = make_problem(cfg, **kw)
    gp = oracle_gp(p)
    X = candidates(p, 5, q=q)
    m1, c1 = gp.posterior(X)  # L^-T GEMM form (fast_pred_var)
    m2, c2 = gp.posterior_direct(X)  # cholesky_solve form
    scale = float(gp.y_std) ** 2 * float(p.hypers.get("outputscale", torch.ones(1)).max() if p.hypers.get("outputscale") is not None else 1.0)
    assert rel_err(m1, m2) < 1e-11
    assert rel_err(c1, c2, scale=scale) < 1e-10
    mq, Sqb, Sqq = gp.posterior_minimal_blocks(X, p.train_X)
    assert rel_err(mq, m2) < 1e-11 and rel_err(Sqq, c2, scale=scale) < 1e-10


@pytest.mark.parametrize("cfg,kw,q", [("C1", {}, 1), ("C1", {}, 3), ("C2", {"n": 40}, 4), ("C4", {"n": 30, "M": 2}, 2)])
def test_cached_equals_uncached_and_minimal(cfg, kw, q):
    """test_logei.py:559-676 (cache_root True == False, value and X-gradient) and test_low_rank.py:45-207."""
    p = make_problem(cfg, **kw)
    acq = oracle_acq(p, oracle_gp(p))
    X = candidates(p, 4, q=q)
    a1, g1 = acq.value_and_grad(X, "faithful")
    a2, g2 = acq.value_and_grad(X, "minimal")
    a3, g3 = acq.value_and_grad(X, "uncached")
    assert rel_err(a2, a1, scale=0.0) < 1e-10 and rel_err(a3, a1, scale=0.0) < 1e-8
    assert rel_err(g2, g1) < 1e-9 and rel_err(g3, g1) < 1e-6


def test_oracle_outputs_match_committed_fixture(cfg_gold):
    """The oracle itself is pinned: any change to oracle/ that moves a number fails here."""
    for tag in ("C1", "C2", "C3", "C4"):
        c = cfg_gold[tag]
        p = make_problem(c["cfg"], **c["kw"])
        gp = oracle_gp(p)
        acq = oracle_acq(p, gp)
        mean, cov = gp.posterior(c["X"])
        val, grad = acq.value_and_grad(c["X"], "faithful")
        assert rel_err(mean, c["mean"]) < 1e-12 and rel_err(cov, c["cov"]) < 1e-10
        assert rel_err(val, c["acq"], scale=0.0) < 1e-10 and rel_err(grad, c["grad"]) < 1e-8
        raw = candidates(p, 128, q=c["q"], seed=c["raw_seed"])
        with torch.no_grad():
            assert int(torch.argmax(acq.forward(raw))) == c["raw_argmax"]


def test_psd_safe_cholesky_semantics():
    from oracle.gp import NanError, NotPSDError, NumericalWarning, psd_safe_cholesky

    A = torch.tensor([[1.0, 1.0], [1.0, 1.0 - 1e-12]], dtype=torch.float64)
    with pytest.warns(NumericalWarning):
        L = psd_safe_cholesky(A)
    assert torch.allclose(L @ L.mT, A, atol=1e-7)
    with pytest.raises(NanError):
        psd_safe_cholesky(torch.full((2, 2), float("nan"), dtype=torch.float64))
    with pytest.raises(NotPSDError), pytest.warns(NumericalWarning):
        psd_safe_cholesky(-torch.eye(2, dtype=torch.float64))


def test_base_sample_layout_of_qlognei():
    """sampling/normal.py:68-131: appending q points keeps the baseline columns of the base samples."""
    zb, zq = qmc.qlognei_base_samples(S=8, n_b=5, q=3, seed=1234)
    assert zb.shape == (8, 5) and zq.shape == (8, 3)
    assert torch.equal(zb, qmc.sobol_normal(5, 8, 1234))
    assert torch.equal(zq, qmc.sobol_normal(8, 8, 1234)[:, 5:])


def test_plain_ei_reduction_known_answers():
    """qEI / qNEI reductions (monte_carlo.py:410,180-252) on hand-made samples: the known answers of
    test/acquisition/test_monte_carlo.py:112-136 (zero samples: 0 at best_f = 0, 1 at best_f = -1) and the
    gradient conventions of torch (amax: even split over ties, clamp_min: passes x >= 0)."""
    from oracle.qei import _plain_reduction

    zeros = torch.zeros(2, 1, 1, dtype=torch.float64)  # S x B x q
    assert float(_plain_reduction(zeros, torch.tensor(0.0))) == 0.0
    assert float(_plain_reduction(zeros, torch.tensor(-1.0))) == 1.0
    obj = torch.tensor([[[0.5, 2.0, 2.0]], [[-1.0, -3.0, -2.0]]], dtype=torch.float64, requires_grad=True)  # S=2, B=1, q=3
    val = _plain_reduction(obj, torch.tensor(0.0))
    assert float(val) == 1.0  # mean(2, 0)
    (g,) = torch.autograd.grad(val.sum(), obj)
    assert torch.equal(g, torch.tensor([[[0.0, 0.25, 0.25]], [[0.0, 0.0, 0.0]]], dtype=torch.float64))


def test_oracle_ei_family_consistency():
    """qNEI oracle == mean_s relu(max_j f_sj - best_f_s) on the qLogNEI sample path; qLogEI ~ log qEI up to the
    smoothing (tau_relu = 1e-6, tau_max = 1e-2) where improvement is common."""
    from oracle.qei import OracleQEI, OracleQLogEI, OracleQNEI

    p = make_problem("C2", n=40)
    ogp = oracle_gp(p)
    X = candidates(p, 4, q=3)
    nei, lognei = OracleQNEI(ogp, p.train_X, p.mc_samples), OracleQLogNEI(ogp, p.train_X, p.mc_samples)
    f = lognei.samples_faithful(X)
    manual = (f.amax(-1) - lognei.best_f.unsqueeze(1)).clamp_min(0).mean(0)
    _close(nei.forward(X), manual)
    with torch.no_grad():
        best_f = ogp.posterior(p.train_X)[0].min()
    ei, logei = OracleQEI(ogp, best_f, p.mc_samples), OracleQLogEI(ogp, best_f, p.mc_samples)
    assert float((logei.forward(X) - ei.forward(X).log()).abs().max()) < 0.05


# ---- exact marginal log likelihood oracle (oracle/mll.py) ---------------------------------------------------------------
def test_mll_oracle_against_torch_distributions():
    """The oracle's log_prob / prior restatements agree with torch.distributions (what gpytorch's MultivariateNormal,
    LogNormalPrior and GammaPrior wrap) and its analytic pieces with autograd."""
    import math

    from oracle import mll as omll

    g = torch.Generator().manual_seed(0)
    n, d = 40, 3
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    y = torch.randn(n, generator=g, dtype=torch.float64)
    ls = torch.tensor([0.4, 1.1, 0.7], dtype=torch.float64)
    noise, mean, os = (torch.tensor(v, dtype=torch.float64) for v in (0.05, 0.3, 1.4))
    for kernel in ("rbf", "matern52"):
        K = omll.train_kernel(X, kernel, ls, os) + noise * torch.eye(n, dtype=torch.float64)
        ref = torch.distributions.MultivariateNormal(mean.expand(n), covariance_matrix=K).log_prob(y)
        got = omll.log_marginal(X, y, kernel, ls, noise, mean, os)
        assert abs(float(got - ref)) < 1e-10 * abs(float(ref))
    x = torch.tensor([0.03, 0.5, 2.0, 40.0], dtype=torch.float64)
    assert torch.allclose(omll.lognormal_log_prob(x, 1.2, math.sqrt(3.0)),
                          torch.distributions.LogNormal(torch.tensor(1.2, dtype=torch.float64), torch.tensor(3.0, dtype=torch.float64).sqrt()).log_prob(x), rtol=1e-13, atol=1e-13)  # fmt: skip
    assert torch.allclose(omll.gamma_log_prob(x, 2.0, 0.15), torch.distributions.Gamma(torch.tensor(2.0, dtype=torch.float64), torch.tensor(0.15, dtype=torch.float64)).log_prob(x), rtol=1e-13, atol=1e-13)
    # the full objective = (log_prob + priors) / n, finite-difference check of the autograd gradient
    h = {"lengthscale": ls, "noise": noise, "mean": mean, "outputscale": None}
    v, grads = omll.value_and_grad(omll.exact_mll, X, y, "rbf", h)
    eps = 1e-6
    hp = dict(h, noise=noise + eps)
    hm = dict(h, noise=noise - eps)
    fd = (omll.exact_mll(X, y, "rbf", hp["lengthscale"], hp["noise"], mean) - omll.exact_mll(X, y, "rbf", hm["lengthscale"], hm["noise"], mean)) / (2 * eps)
    assert abs(float(fd - grads["noise"])) < 1e-6 * max(abs(float(fd)), 1.0)


def test_fit_host_prior_terms_match_the_oracle():
    """botorch_b200.fit's closed-form prior value/derivative (host-side plumbing) == oracle restatement + autograd."""
    from botorch_b200 import fit as bfit
    from oracle import mll as omll

    x = torch.tensor([0.03, 0.5, 2.0, 40.0], dtype=torch.float64, requires_grad=True)
    for (fn, ofn, args) in ((bfit._lognormal, omll.lognormal_log_prob, (-4.0, 1.0)), (bfit._gamma, omll.gamma_log_prob, (2.0, 0.15))):
        lp, dlp = fn(x.detach(), *args)
        ref = ofn(x, *args)
        (gref,) = torch.autograd.grad(ref.sum(), x)
        assert torch.allclose(lp, ref.detach(), rtol=1e-14, atol=1e-14)
        assert torch.allclose(dlp, gref, rtol=1e-13, atol=1e-13)
    assert bfit.lengthscale_prior_params(20) == (2.0**0.5 + 0.5 * __import__("math").log(20), 3.0**0.5)


# ---- pathwise (Matheron) sampling oracle (oracle/pathwise.py) ----------------------------------------------------------
def test_pathwise_oracle_is_consistent_with_the_exact_posterior():
    """No reference fixture exists for sample paths; the restatement is pinned by its defining properties against the
    posterior oracle: zero draws give the posterior mean, and the path ensemble reproduces the posterior moments (the
    covariance up to the random-feature approximation of the prior)."""
    from oracle import pathwise as opw
    from oracle.gp import GPHyper, OracleGP

    g = torch.Generator().manual_seed(0)
    n, d, F, S = 30, 2, 4096, 3000
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    y = torch.sin(4 * X[:, :1]) + 0.1 * torch.randn(n, 1, generator=g, dtype=torch.float64)
    h = GPHyper("rbf", torch.tensor([0.4, 0.7], dtype=torch.float64), torch.tensor(0.05, dtype=torch.float64),
                torch.tensor(0.1, dtype=torch.float64), None)
    gp = OracleGP(X, y, h, standardize=True)
    Xc = torch.rand(12, d, generator=g, dtype=torch.float64)
    mean, cov = gp.posterior(Xc)
    args = (Xc, X, gp.train_targets, "rbf", h.lengthscale, h.noise, h.mean, None)
    W = torch.randn(F // 2, d, generator=g, dtype=torch.float64)
    zero = opw.matheron_paths(*args, W, torch.zeros(1, F, dtype=torch.float64), torch.zeros(1, n, dtype=torch.float64),
                              float(gp.y_mean), float(gp.y_std))
    assert torch.allclose(zero[0], mean.reshape(-1), rtol=1e-10, atol=1e-12)
    w = torch.randn(S, F, generator=g, dtype=torch.float64)
    eps = torch.randn(S, n, generator=g, dtype=torch.float64)
    vals = opw.matheron_paths(*args, W, w, eps, float(gp.y_mean), float(gp.y_std))
    sd = cov.diagonal(dim1=-2, dim2=-1).reshape(-1).sqrt()
    assert float(((vals.mean(0) - mean.reshape(-1)).abs() / sd).max()) < 0.12
    emp = torch.cov(vals.T)
    assert float((emp - cov.reshape(12, 12)).abs().max() / cov.abs().max()) < 0.15


# ---- outcome constraints: feasibility weights of the product (botorch_b200/constrained.py) and of the
# oracle (oracle/constrained.py) against literals minted from the reference's utils/safe_math.py + utils/objective.py
_FEAS_X = [-30.0, -2.5, -0.3, 0.0, 0.4, 3.0, 50.0]
_LOG_FATMOID = [-7.247053179813573, -2.7539872207593805, -0.976299726026775, -0.6931471805599453, -0.4169842841295436,
                -0.04952450151883789, -0.0002605454114726355]  # fmt: skip
_LOG_FATMOID_TAU_HALF = [-8.613579709661215, -3.874534075858838, -1.2751478288557359, -0.6931471805599453,
                         -0.2615149935538013, -0.015176570258031662, -6.58991390632818e-05]  # fmt: skip
_LOGEXPIT = [-30.000000000000092, -2.5788897342925496, -0.8543552444685272, -0.6931471805599453, -0.5130152523999526,
             -0.04858735157374206, -1.9287498479639178e-22]  # fmt: skip


def _feasibility_impls():
    from botorch_b200 import constrained as prod
    from oracle import constrained as orc

    return [prod, orc]


def test_feasibility_weights_match_reference_literals():
    x = torch.tensor(_FEAS_X, dtype=torch.float64)
    for impl in _feasibility_impls():
        assert rel_err(impl.log_fatmoid(x), torch.tensor(_LOG_FATMOID, dtype=torch.float64), scale=1.0) < 1e-14
        assert rel_err(impl.log_fatmoid(x, tau=0.5), torch.tensor(_LOG_FATMOID_TAU_HALF, dtype=torch.float64), scale=1.0) < 1e-14
        got, want = impl.logexpit(x), torch.tensor(_LOGEXPIT, dtype=torch.float64)
        assert bool(((got - want).abs() <= 1e-14 * want.abs().clamp_min(1e-300) + 1e-36).all())
        assert rel_err(impl.fatmoid(x).log(), impl.log_fatmoid(x), scale=1.0) < 1e-13


def test_smoothed_feasibility_indicator_semantics():
    """utils/objective.py:131-185: product over constraints of sigmoid(-c/eta) (sum of logs), one eta per constraint;
    the hard indicator is c <= 0 for every constraint."""
    from botorch_b200 import constrained as prod
    from oracle import constrained as orc

    g = torch.Generator().manual_seed(0)
    Z = torch.randn(5, 3, 4, 2, generator=g, dtype=torch.float64)
    cons = [lambda s: s[..., 1] - 0.2, lambda s: 0.5 * s[..., 0] + s[..., 1]]
    eta = torch.tensor([1e-2, 0.3], dtype=torch.float64)
    for fat in (True, False):
        f = prod.log_fatmoid if fat else prod.logexpit
        want = f(-cons[0](Z) / eta[0]) + f(-cons[1](Z) / eta[1])
        got = prod.compute_smoothed_feasibility_indicator(cons, Z, eta, log=True, fat=fat)
        assert got.shape == Z.shape[:-1] and torch.equal(got, want)
        assert rel_err(orc.smoothed_log_feasibility(cons, Z, eta, fat), want, scale=1.0) < 1e-15
        lin = prod.compute_smoothed_feasibility_indicator(cons, Z, eta, log=False, fat=fat)
        assert rel_err(lin, want.exp(), scale=1.0) < 1e-14
    hard = (cons[0](Z) <= 0) & (cons[1](Z) <= 0)
    assert torch.equal(prod.compute_feasibility_indicator(cons, Z), hard)
    assert torch.equal(orc.feasibility(cons, Z), hard)
    assert torch.equal(prod.compute_feasibility_indicator(None, Z), torch.ones(5, 3, 4, dtype=torch.bool))
    with pytest.raises(ValueError):
        prod.compute_smoothed_feasibility_indicator(cons, Z, torch.ones(3), log=True)
    # a scalar eta is broadcast as torch.full((n,), eta) in the default dtype, exactly as the reference does (:167-168)
    a = prod.compute_smoothed_feasibility_indicator(cons, Z, 0.1, log=True)
    b = prod.compute_smoothed_feasibility_indicator(cons, Z, torch.full((2,), 0.1), log=True)
    assert torch.equal(a, b) and a.dtype == torch.float64


@pytest.mark.skipif(not refstub.available(), reason="/root/reference is not present on this machine")
def test_feasibility_weights_match_the_live_reference():
    refstub.load()
    from botorch.utils.objective import compute_feasibility_indicator, compute_smoothed_feasibility_indicator

    from botorch_b200 import constrained as prod

    g = torch.Generator().manual_seed(1)
    Z = torch.randn(7, 2, 3, 3, generator=g, dtype=torch.float64)
    cons = [lambda s: s[..., 1] - 0.2, lambda s: s[..., 2] * s[..., 0] - 0.1]
    for eta in (1e-3, torch.tensor([1e-2, 0.5], dtype=torch.float64)):
        for fat in (True, False):
            want = compute_smoothed_feasibility_indicator(cons, Z, eta, log=True, fat=fat)
            got = prod.compute_smoothed_feasibility_indicator(cons, Z, eta, log=True, fat=fat)
            assert rel_err(got, want, scale=1.0) < 1e-13
    assert torch.equal(prod.compute_feasibility_indicator(cons, Z), compute_feasibility_indicator(cons, Z))


def test_constrained_oracle_reduces_to_the_plain_oracle_without_constraints():
    """One outcome, no constraints: oracle/constrained.py must reproduce oracle/qlognei.py (pinned against the reference's
    golden vectors above) -- same Sobol dimensions, same incumbent, same reductions."""
    from oracle.constrained import OracleConstrainedQLogNEI

    p = make_problem("C1")
    gp = oracle_gp(p)
    plain = OracleQLogNEI(gp, p.train_X, num_samples=32, seed=5)
    split = OracleConstrainedQLogNEI([gp], p.train_X, [], 1e-3, obj_idx=0, obj_weight=1.0, num_samples=32, seed=5)
    X = candidates(p, 4, 2, seed=1)
    assert rel_err(split.forward(X), plain.forward(X), scale=1.0) < 1e-12
    v1, g1 = split.value_and_grad(X)
    v2, g2 = plain.value_and_grad(X)
    assert rel_err(g1, g2, scale=float(g2.abs().max())) < 1e-11


# ---- the GP algebra against 50-digit mpmath vectors (independent of torch: tests/golden/make_mpmath_golden.py) -------
def _mp_cases():
    from tests.util import mpmath_cases

    return mpmath_cases()


@pytest.mark.parametrize("case,p", _mp_cases(), ids=lambda v: v["name"] if isinstance(v, dict) else "")
def test_oracle_gp_algebra_matches_mpmath(case, p):
    """oracle/gp.py + oracle/qlognei.py (the restatement of gpytorch 1.12 / linear_operator 0.5.2 behind
    models/gpytorch.py:446, utils/low_rank.py:118-162, acquisition/logei.py:391-481) against the published formulas
    evaluated in 50-digit arithmetic without torch.  This is the pin SURVEY.md 8(c) says the reference cannot give."""
    from tests.util import oracle_acq, oracle_gp

    t = lambda v: torch.tensor(v, dtype=torch.float64)  # noqa: E731
    ogp = oracle_gp(p)
    assert abs(float(ogp.y_mean) - case["y_mean"]) <= 1e-14 * max(1.0, abs(case["y_mean"]))
    assert abs(float(ogp.y_std) - case["y_std"]) <= 1e-14 * case["y_std"]
    X = t(case["X"])  # B x q x d
    prior = float(ogp.y_std**2 * (case["outputscale"] or 1.0))
    for form in (ogp.posterior, ogp.posterior_direct):
        mean, cov = form(X)
        assert rel_err(mean, t(case["post_mean"])) < 1e-13  # measured over the 48 cases: 1.2e-15
        assert rel_err(cov, t(case["post_cov"]), scale=prior) < 1e-13  # measured 8e-16;  # cancellation k** - |L^-1 k|^2: prior-scale form (H1)
    oacq = oracle_acq(p, ogp, seed=case["seed"])
    assert torch.equal(oacq.Z_b, t(case["Z_b"])) and torch.equal(oacq.Z_q(p.q), t(case["Z_q"]))
    assert rel_err(oacq.baseline_L, t(case["baseline_L"]), scale=prior**0.5) < 1e-12  # measured 6e-14
    assert rel_err(oacq.best_f, t(case["best_f"])) < 1e-12  # measured 4e-14
    ref_samples = t(case["samples"]).permute(1, 0, 2)  # S x B x q
    for sampler in (oacq.samples_faithful, oacq.samples_minimal):
        assert rel_err(sampler(X), ref_samples) < 1e-11  # measured 1.1e-13
    for mode in ("faithful", "minimal"):
        val, grad = oacq.value_and_grad(X, mode=mode)
        assert rel_err(val, t(case["acq"]), scale=0.0) < 1e-11, mode  # measured 9e-14
        gref = t(case["grad"])
        gerr = ((grad - gref).flatten(1).abs().max(1).values / gref.flatten(1).abs().max(1).values).max()
        assert float(gerr) < 1e-10, mode  # measured 1.2e-12


def test_oracle_nofat_reductions_match_reference_fixtures():
    """fat = False (acquisition/logei.py:122,532): the oracle's log_softplus / smooth_amax and the per-sample chain against
    fixtures minted from the reference's own utils/safe_math.py (tests/golden/make_golden.py::_nofat_part)."""
    import os

    from oracle import safe_math as osm

    fx = torch.load(os.path.join(os.path.dirname(__file__), "golden", "reference_nofat.pt"))
    for name, c in fx.items():
        x = c.get("x")
        if name.startswith("log_softplus_tau"):
            tau = float(name.split("tau")[1])
            xg = x.clone().requires_grad_(True)
            v = osm.log_softplus(xg, tau=tau)
            (g,) = torch.autograd.grad(v.sum(), xg)
            assert torch.equal(v.detach(), c["value"]) and torch.equal(g, c["grad"]), name
        elif name.startswith("smooth_amax_tau"):
            tau = float(name.split("tau")[1])
            xg = x.clone().requires_grad_(True)
            v = osm.smooth_amax(xg, dim=-1, tau=tau)
            (g,) = torch.autograd.grad(v.sum(), xg)
            assert torch.allclose(v.detach(), c["value"], rtol=1e-15, atol=0) and torch.allclose(g, c["grad"], rtol=1e-14, atol=1e-300), name
        elif name == "smooth_amax_inf":
            v = osm.smooth_amax(x, dim=-1, tau=1e-2)
            assert torch.equal(torch.isinf(v), torch.isinf(c["value"])) and torch.equal(v[torch.isinf(v)], c["value"][torch.isinf(v)])
            assert torch.allclose(v[~torch.isinf(v)], c["value"][~torch.isinf(v)], rtol=1e-15)
        elif name == "qlognei_chain_nofat":
            og = c["obj"].clone().requires_grad_(True)
            acq = osm.qlognei_reduction(og, c["best_f"].squeeze(-1).unsqueeze(-1) if False else c["best_f"], 1e-6, 1e-2, fat=False)
            (g,) = torch.autograd.grad(acq.sum(), og)
            assert torch.allclose(acq.detach(), c["value"], rtol=1e-14) and torch.allclose(g, c["grad"], rtol=1e-12, atol=1e-300)
