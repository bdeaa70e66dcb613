"""GPU parity of the rest of the improvement family on the same fused kernels: qNEI, qLogEI, qEI
values and X-gradients against the CPU oracle (oracle/qei.py), through the public classes and torch.autograd."""
import warnings

import pytest
import torch

import botorch_b200 as bb
from botorch_b200.synthetic import make_problem
from oracle.qei import OracleQEI, OracleQLogEI, OracleQNEI
from tests.test_gpu_api import _model
from tests.util import candidates, oracle_gp, rel_err

pytestmark = pytest.mark.gpu
RTOL = 1e-9
CASES = [("C1", {}, 1, 10), ("C2", {}, 4, 16), ("C3", {"n": 256}, 8, 6), ("C4", {"M": 5}, 4, 7),
         ("C1", {"n": 150, "noise": 1e-4}, 3, 6)]  # densely sampled, noise floor: the root must not square cond(K)
# the noisy incumbent (max over n baseline draws) is out of reach of the plain qNEI at the bigger training sets: it is
# identically 0 there -- the reason qLogNEI exists -- so its parity is checked where improvement is possible
NEI_CASES = [("C1", {}, 1, 10), ("C2", {"n": 30}, 4, 16), ("C3", {"n": 100}, 8, 6), ("C4", {"M": 5}, 4, 7)]


def _sampler(p):
    return bb.SobolQMCNormalSampler(sample_shape=torch.Size([p.mc_samples]), seed=1234)


def _value_and_grad(acqf, X, dev):
    Xd = X.to(dev).requires_grad_(True)
    val = acqf(Xd)
    (g,) = torch.autograd.grad(val.sum(), Xd)
    acqf.check_numerics()
    return val.detach(), g


def _near_best(p, B, q):
    """t-batches around the best observed points: random candidates almost never improve on the noisy incumbent, which
    would make the plain (non-log) parity check compare zeros."""
    order = torch.argsort(p.train_Y.view(-1), descending=True)
    top = order[torch.arange(B * q) % min(8, order.numel())]
    gen = torch.Generator().manual_seed(5)
    X = p.train_X[top] + 0.05 * torch.randn(B * q, p.d, generator=gen, dtype=torch.float64)
    return X.clamp(0.0, 1.0).reshape(B, q, p.d)


def _incumbent(ogp, X):
    """A best_f that leaves a healthy share of the samples improving: the median over the t-batches of the largest
    posterior mean within each q-batch (a scalar, or one value per model)."""
    with torch.no_grad():
        mean, _ = ogp.posterior(X)
    return mean.amax(dim=-1).median(dim=0).values


@pytest.mark.parametrize("cfg,kw,q,B", NEI_CASES)
def test_qnei_matches_oracle(cuda, cfg, kw, q, B):
    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    oacq = OracleQNEI(ogp, p.train_X, num_samples=p.mc_samples, seed=1234)
    X = torch.cat([_near_best(p, B - 2, q), candidates(p, 2, q=q)])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", bb.exceptions.NumericsWarning)
        acqf = bb.qNoisyExpectedImprovement(_model(p, cuda), p.train_X.to(cuda), sampler=_sampler(p), prune_baseline=False)
    val, g = _value_and_grad(acqf, X, cuda)
    ref, g_ref = oacq.value_and_grad(X)
    assert val.shape == (B,) and int((ref > 0).sum()) >= 3
    assert rel_err(val, ref) < RTOL
    assert rel_err(g, g_ref) < 1e-8


@pytest.mark.parametrize("cfg,kw,q,B", CASES)
@pytest.mark.parametrize("log", [True, False])
def test_qei_and_qlogei_match_oracle(cuda, cfg, kw, q, B, log):
    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    X = candidates(p, B, q=q)
    best_f = _incumbent(ogp, X)
    model = _model(p, cuda)
    if log:
        oacq = OracleQLogEI(ogp, best_f, num_samples=p.mc_samples, seed=1234)
        acqf = bb.qLogExpectedImprovement(model, best_f.to(cuda), sampler=_sampler(p))
    else:
        oacq = OracleQEI(ogp, best_f, num_samples=p.mc_samples, seed=1234)
        with pytest.warns(bb.exceptions.NumericsWarning):
            acqf = bb.qExpectedImprovement(model, best_f.to(cuda), sampler=_sampler(p))
    val, g = _value_and_grad(acqf, X, cuda)
    ref, g_ref = oacq.value_and_grad(X)
    assert val.shape == (B,) and (log or int((ref > 0).sum()) >= B // 2)
    assert rel_err(val, ref, scale=None if not log else 0.0) < RTOL
    assert rel_err(g, g_ref) < 1e-8
    # no-grad evaluation (raw-sample phase) gives the same numbers
    with torch.no_grad():
        val2 = acqf(X.to(cuda))
    assert torch.equal(val2, val)


def test_qei_known_answers(cuda):
    """test/acquisition/test_monte_carlo.py:112-136 in spirit: with the incumbent far below / above every sample the
    plain qEI equals mean_s max_j f_sj - best_f / exactly 0 with a zero gradient."""
    p = make_problem("C2", n=40)
    ogp = oracle_gp(p)
    model = _model(p, cuda)
    X = candidates(p, 3, q=2)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", bb.exceptions.NumericsWarning)
        hi = bb.qExpectedImprovement(model, 1e6, sampler=_sampler(p))
        lo = bb.qExpectedImprovement(model, -1e6, sampler=_sampler(p))
    val, g = _value_and_grad(hi, X, cuda)
    assert float(val.abs().max()) == 0.0 and float(g.abs().max()) == 0.0
    val, _ = _value_and_grad(lo, X, cuda)
    samples = OracleQEI(ogp, -1e6, num_samples=p.mc_samples, seed=1234).samples(X)
    assert rel_err(val, samples.amax(-1).mean(0) + 1e6) < RTOL


def test_qnei_prunes_baseline_by_default(cuda):
    """prune_baseline=True (the reference default for qNEI) keeps only baseline points that are the best one in some
    joint posterior draw; the pruned set is a subset of train_X and the value is finite and differentiable."""
    p = make_problem("C2", n=60)
    model = _model(p, cuda)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", bb.exceptions.NumericsWarning)
        acqf = bb.qNoisyExpectedImprovement(model, p.train_X.to(cuda), sampler=_sampler(p))
    Xb = acqf.X_baseline.cpu()
    assert 1 <= Xb.shape[0] < p.train_X.shape[0]
    assert all(bool((p.train_X == row).all(dim=-1).any()) for row in Xb)
    # value on the pruned baseline == oracle with the same baseline
    ogp = oracle_gp(p)
    oacq = OracleQNEI(ogp, Xb, num_samples=p.mc_samples, seed=1234)
    X = torch.cat([_near_best(p, 4, 3), candidates(p, 1, q=3)])
    val, g = _value_and_grad(acqf, X, cuda)
    ref, g_ref = oacq.value_and_grad(X)
    assert float(ref.max()) > 0.0
    assert rel_err(val, ref) < RTOL
    assert rel_err(g, g_ref) < 1e-8


@pytest.mark.parametrize("kind", ["qlognei", "qnei", "qlogei", "qei"])
def test_linear_objective_weight(cuda, kind):
    """objective=LinearMCObjective(weights=[w]) on the one-outcome model (objective.py:217-256; w = -1 is how the
    reference minimises): samples are scaled by w before the incumbent and the improvement are formed."""
    w = -1.0 if kind != "qlogei" else -0.5
    p = make_problem("C2", n=30)
    ogp = oracle_gp(p)
    model = _model(p, cuda)
    objective = bb.LinearMCObjective(torch.tensor([w], dtype=torch.float64, device=cuda))
    order = torch.argsort(p.train_Y.view(-1))  # candidates near the SMALLEST observations: w < 0 rewards those
    gen = torch.Generator().manual_seed(11)
    X = (p.train_X[order[torch.arange(12) % 6]] + 0.05 * torch.randn(12, p.d, generator=gen, dtype=torch.float64))
    X = torch.cat([X.clamp(0, 1).reshape(4, 3, p.d), candidates(p, 2, q=3)])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", bb.exceptions.NumericsWarning)
        if kind in ("qlognei", "qnei"):
            ocls = OracleQNEI if kind == "qnei" else __import__("oracle.qlognei", fromlist=["x"]).OracleQLogNEI
            oacq = ocls(ogp, p.train_X, num_samples=p.mc_samples, seed=1234, obj_weight=w)
            cls = bb.qNoisyExpectedImprovement if kind == "qnei" else bb.qLogNoisyExpectedImprovement
            acqf = cls(model, p.train_X.to(cuda), sampler=_sampler(p), objective=objective, prune_baseline=False)
        else:
            with torch.no_grad():
                mean, _ = ogp.posterior(X)
            best_f = (w * mean).amax(dim=-1).median(dim=0).values
            ocls = OracleQEI if kind == "qei" else OracleQLogEI
            oacq = ocls(ogp, best_f, num_samples=p.mc_samples, seed=1234, obj_weight=w)
            cls = bb.qExpectedImprovement if kind == "qei" else bb.qLogExpectedImprovement
            acqf = cls(model, best_f.to(cuda), sampler=_sampler(p), objective=objective)
    val, g = _value_and_grad(acqf, X, cuda)
    ref, g_ref = oacq.value_and_grad(X)
    log = kind.startswith("qlog")
    assert log or int((ref > 0).sum()) >= 2
    assert rel_err(val, ref, scale=0.0 if log else None) < RTOL
    assert rel_err(g, g_ref) < 1e-8
    # the weight matters: the identity objective gives different numbers
    if kind == "qlognei":
        plain = bb.qLogNoisyExpectedImprovement(model, p.train_X.to(cuda), sampler=_sampler(p), prune_baseline=False)
        with torch.no_grad():
            assert not torch.allclose(plain(X.to(cuda)), val)


def test_prune_inferior_points_with_objective(cuda):
    """utils.py:252-341 with a negating objective keeps the points that are best under w * f."""
    p = make_problem("C2", n=40)
    model = _model(p, cuda)
    obj = bb.LinearMCObjective(torch.tensor([-1.0], dtype=torch.float64, device=cuda))
    keep_max = bb.prune_inferior_points(model, p.train_X.to(cuda))
    keep_min = bb.prune_inferior_points(model, p.train_X.to(cuda), objective=obj)
    y = p.train_Y.view(-1)
    def _rows(K):
        return [int((p.train_X == r).all(-1).nonzero()[0, 0]) for r in K.cpu()]
    assert int(torch.argmax(y)) in _rows(keep_max) and int(torch.argmin(y)) in _rows(keep_min)
    assert set(_rows(keep_max)) != set(_rows(keep_min))


@pytest.mark.parametrize("cfg,kw,q,B", [("C2", {}, 4, 16), ("C1", {}, 1, 8), ("C3", {"n": 300}, 8, 6), ("C4", {"M": 4}, 3, 5)])
def test_fat_false_uses_log_softplus_and_smooth_amax(cuda, cfg, kw, q, B):
    """acquisition/logei.py:122,532: fat = False switches log_fatplus -> log_softplus and fatmax -> smooth_amax (both pinned
    to the reference's utils/safe_math.py in tests/test_oracle_golden.py); qLogNEI and qLogEI, values and gradients."""
    import botorch_b200 as bb
    from botorch_b200.synthetic import make_problem
    from oracle.qlognei import OracleQLogNEI
    from tests.test_gpu_qlognei import _grad_rel
    from tests.util import candidates, oracle_gp, rel_err

    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    oacq = OracleQLogNEI(ogp, p.train_X, num_samples=p.mc_samples, seed=1234, fat=False)
    h = p.hypers
    if p.kernel == "rbf":
        model = bb.SingleTaskGP(p.train_X.to(cuda), p.train_Y.to(cuda), lengthscale=h["lengthscale"], noise=h["noise"],
                                mean_constant=h["mean"])  # fmt: skip
    else:
        model = bb.SaasFullyBayesianSingleTaskGP(p.train_X.to(cuda), p.train_Y.to(cuda))
        model.load_mcmc_samples(h)
    acqf = bb.qLogNoisyExpectedImprovement(model, p.train_X.to(cuda), fat=False,
                                           sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234))  # fmt: skip
    X = candidates(p, B, q=q)
    Xd = X.to(cuda).requires_grad_(True)
    val = acqf(Xd)
    (grad,) = torch.autograd.grad(val.sum(), Xd)
    ref_val, ref_grad = oacq.value_and_grad(X, mode="faithful")
    assert rel_err(val, ref_val, scale=0.0) < 1e-9
    assert _grad_rel(grad, ref_grad) < 1e-9
    fat_val = bb.qLogNoisyExpectedImprovement(model, p.train_X.to(cuda),
                                              sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234))(X.to(cuda))  # fmt: skip
    assert float((fat_val.cpu() - ref_val).abs().max()) > 1e-6  # the two smoothings really differ
