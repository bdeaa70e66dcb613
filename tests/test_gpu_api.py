"""GPU parity of the reference-facing Python API (Model.posterior, posterior sampling, qLogNEI forward/backward via
torch.autograd, optimize_acqf plumbing, Thompson sampling) against the CPU oracle."""
import math
import warnings

import pytest
import torch

import botorch_b200 as bb
from botorch_b200.synthetic import make_problem
from oracle import qmc
from tests.util import candidates, oracle_acq, oracle_gp, rel_err

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def _model(p, dev):
    h = p.hypers
    if p.kernel == "rbf":
        return bb.SingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev), lengthscale=h["lengthscale"], noise=h["noise"],
                               mean_constant=h["mean"], outcome_transform="DEFAULT" if p.standardize else None)  # fmt: skip
    m = bb.SaasFullyBayesianSingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev))
    m.load_mcmc_samples(h)
    return m


def _acqf(p, model, dev, **kw):
    sampler = bb.SobolQMCNormalSampler(sample_shape=torch.Size([p.mc_samples]), seed=1234)
    return bb.qLogNoisyExpectedImprovement(model, p.train_X.to(dev), sampler=sampler, **kw)


@pytest.mark.parametrize("cfg,kw,bshape,q", [("C1", {}, (5,), 1), ("C2", {}, (3, 2), 4), ("C2", {"n": 77}, (), 3)])
def test_posterior_mean_covariance_and_grad(cuda, cfg, kw, bshape, q):
    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    model = _model(p, cuda)
    X = candidates(p, max(math.prod(bshape), 1), q=q).reshape(*bshape, q, p.d)
    Xd = X.to(cuda).requires_grad_(True)
    post = model.posterior(Xd)
    Xo = X.clone().requires_grad_(True)
    m_ref, c_ref = ogp.posterior_direct(Xo)
    assert post.mean.shape == (*bshape, q, 1) and post.variance.shape == (*bshape, q, 1)
    assert rel_err(post.mean.squeeze(-1), m_ref) < RTOL
    # variance tolerance relative to the prior scale: cancellation k** - |L^-1 k|^2
    prior = float(ogp.y_std) ** 2
    assert rel_err(post.distribution.covariance_matrix, c_ref, scale=prior) < RTOL
    # gradient of a scalar functional of mean and covariance
    w = torch.linspace(0.5, 1.5, q, dtype=torch.float64)
    f = (post.mean.squeeze(-1) * w.to(cuda)).sum() + (post.distribution.covariance_matrix**2).sum()
    (g,) = torch.autograd.grad(f, Xd)
    f_ref = (m_ref * w).sum() + (c_ref**2).sum()
    (g_ref,) = torch.autograd.grad(f_ref, Xo)
    assert rel_err(g, g_ref) < 1e-8


@pytest.mark.parametrize("cfg,kw,q", [("C2", {"n": 60}, 60), ("C3", {"n": 256}, 131), ("C4", {"M": 3}, 45)])
def test_posterior_of_a_large_joint_set(cuda, cfg, kw, q):
    """q beyond the per-candidate kernels' limit (prune_inferior_points evaluates the joint posterior of all of
    X_baseline): the GEMM-shaped path gives the same mean / covariance, and sampling from it works."""
    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    model = _model(p, cuda)
    X = torch.cat([p.train_X[: q // 2], candidates(p, q - q // 2, q=1).reshape(-1, p.d)])
    with torch.no_grad():
        post = model.posterior(X.to(cuda))
    m_ref, c_ref = ogp.posterior_direct(X)
    prior = float(ogp.y_std) ** 2 * (float(p.hypers["outputscale"].max()) if p.hypers.get("outputscale") is not None else 1.0)
    assert rel_err(post.mean.squeeze(-1), m_ref) < RTOL
    assert rel_err(post.distribution.covariance_matrix, c_ref, scale=prior) < RTOL
    sampler = bb.SobolQMCNormalSampler(sample_shape=torch.Size([16]), seed=3)
    samples = sampler(post)
    assert samples.shape[0] == 16 and samples.shape[-2:] == (q, 1) and bool(torch.isfinite(samples).all())
    with pytest.raises(Exception):
        model.posterior(X.to(cuda).requires_grad_(True))


def test_posterior_observation_noise_and_errors(cuda):
    p = make_problem("C1")
    model = _model(p, cuda)
    X = candidates(p, 4, q=2).to(cuda)
    a, b = model.posterior(X), model.posterior(X, observation_noise=True)
    ys2 = float(model.outcome_transform.stdvs) ** 2
    assert torch.allclose(b.variance - a.variance, torch.full_like(a.variance, ys2 * float(p.hypers["noise"])))
    with pytest.raises(bb.exceptions.BotorchTensorDimensionError):
        model.posterior(torch.zeros(3, 5, device=cuda, dtype=torch.float64))
    with pytest.raises(bb.exceptions.BotorchTensorDimensionError):
        bb.SingleTaskGP(p.train_X, p.train_Y.squeeze(-1))
    with pytest.warns(bb.exceptions.InputDataWarning):
        bb.SingleTaskGP(p.train_X.float(), p.train_Y.float())


def test_rsample_from_base_samples_is_mean_plus_chol_z(cuda):
    """test/posteriors/test_gpytorch.py:28-96: samples == mean + chol(Sigma) z."""
    p = make_problem("C2", n=50)
    model = _model(p, cuda)
    X = candidates(p, 6, q=5).to(cuda)
    post = model.posterior(X)
    S = 16
    sampler = bb.SobolQMCNormalSampler(sample_shape=torch.Size([S]), seed=3)
    samples = sampler(post)
    assert samples.shape == (S, 6, 5, 1)
    assert sampler.base_samples.shape == (S, 1, 5)  # collapsed over the t-batch
    L = torch.linalg.cholesky(post.distribution.covariance_matrix.cpu())
    z = sampler.base_samples.cpu().expand(S, 6, 5)
    ref = post.mean.cpu().squeeze(-1) + torch.einsum("bij,sbj->sbi", L, z)
    assert rel_err(samples.squeeze(-1), ref) < 1e-9
    with pytest.raises(RuntimeError):
        post.rsample_from_base_samples(torch.Size([S + 1]), sampler.base_samples.expand(S, 6, 5))
    assert post.rsample(torch.Size([3])).shape == (3, 6, 5, 1)


@pytest.mark.parametrize("cfg,kw", [("C2", {"n": 50}), ("C4", {"M": 3})])
def test_rsample_is_differentiable_through_the_cholesky(cuda, cfg, kw):
    """Reparameterised samples carry gradients w.r.t. X (custom MC acquisition functions call
    sampler(posterior) under autograd, acquisition/monte_carlo.py:116-125): compared with torch autograd through
    mean + cholesky(cov) z on the oracle posterior."""
    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    model = _model(p, cuda)
    X = candidates(p, 4, q=3)
    Xd = X.to(cuda).requires_grad_(True)
    sampler = bb.SobolQMCNormalSampler(sample_shape=torch.Size([8]), seed=5)
    samples = sampler(model.posterior(Xd)).squeeze(-1)  # 8 x 4 x (M) x 3
    w = torch.linspace(0.5, 2.0, 3, dtype=torch.float64)
    (g,) = torch.autograd.grad((samples**2 * w.to(cuda)).sum() + samples.sum(), Xd)
    Xo = X.clone().requires_grad_(True)
    m_ref, c_ref = ogp.posterior(Xo)
    L = torch.linalg.cholesky(c_ref)
    z = sampler.base_samples.cpu().reshape(8, *([1] * (m_ref.dim() - 1)), 3)
    s_ref = m_ref + torch.einsum("...ij,s...j->s...i", L, z.expand(8, *m_ref.shape))
    assert rel_err(samples, s_ref) < 1e-9
    (g_ref,) = torch.autograd.grad((s_ref**2 * w).sum() + s_ref.sum(), Xo)
    assert rel_err(g, g_ref) < 1e-8


@pytest.mark.parametrize("cfg,kw,q,B", [("C1", {}, 1, 10), ("C2", {}, 4, 16), ("C3", {"n": 256}, 8, 6), ("C4", {"M": 5}, 4, 7)])
def test_qlognei_forward_and_autograd_match_oracle(cuda, cfg, kw, q, B):
    p = make_problem(cfg, **kw)
    oacq = oracle_acq(p, oracle_gp(p))
    model = _model(p, cuda)
    acqf = _acqf(p, model, cuda)
    X = candidates(p, B, q=q)
    Xd = X.to(cuda).requires_grad_(True)
    val = acqf(Xd)
    assert val.shape == (B,)
    (g,) = torch.autograd.grad((-val).sum(), Xd)  # the loss gen_candidates_scipy builds
    a_ref, g_ref = oacq.value_and_grad(X, mode="faithful")
    assert rel_err(val, a_ref, scale=0.0) < RTOL
    scale = g_ref.flatten(1).abs().max(dim=1).values
    assert float(((-g.cpu() - g_ref).flatten(1).abs().max(dim=1).values / scale).max()) < RTOL
    # reference-named state
    assert acqf._baseline_L.shape[-2:] == (p.n, p.n) and acqf.q_in == q
    assert rel_err(acqf._baseline_best_f, oacq.best_f) < RTOL
    assert rel_err(acqf.baseline_samples.squeeze(-1), oacq.baseline_samples) < 1e-8
    with torch.no_grad():
        assert torch.equal(acqf(X.to(cuda)), val.detach())
    assert acqf(X[0].to(cuda)).shape == (1,)  # q x d input gains a t-batch dimension


def test_qlognei_pending_points_and_errors(cuda):
    p = make_problem("C2", n=60)
    oacq = oracle_acq(p, oracle_gp(p))
    model = _model(p, cuda)
    X = candidates(p, 5, q=3)
    pend = candidates(p, 1, q=1, seed=9)[0]
    acqf = _acqf(p, model, cuda, X_pending=pend.to(cuda))
    val = acqf(X.to(cuda))
    ref = oacq.forward(torch.cat([X, pend.expand(5, 1, p.d)], dim=-2))
    assert rel_err(val, ref, scale=0.0) < RTOL
    with pytest.raises(ValueError):
        acqf(torch.zeros(p.d, device=cuda, dtype=torch.float64))
    with pytest.raises(bb.exceptions.UnsupportedError):
        _acqf(p, model, cuda, objective=lambda s, X: s)
    with pytest.raises(Exception):  # no CPU path
        acqf(X)


def test_jitter_warning_is_reported_by_check_numerics(cuda):
    """A candidate on a training point makes the q x q Schur block singular: the jitter schedule of psd_safe_cholesky
    runs in the kernel and the host reports it as a NumericalWarning -- one call late, or on check_numerics()."""
    p = make_problem("C1")
    model = _model(p, cuda)
    acqf = _acqf(p, model, cuda)
    X = torch.stack([p.train_X[:2], p.train_X[2:4]]).to(cuda)  # 2 t-batches of q = 2 training points
    with warnings.catch_warnings(record=True) as rec:
        warnings.simplefilter("always")
        val = acqf(X)
        acqf.check_numerics()
    assert bool(torch.isfinite(val).all())
    kinds = {w.category for w in rec}
    _, info, _ = acqf._nei.forward(X.contiguous(), acqf._Zq, acqf.tau_relu, acqf.tau_max, with_grad=False)
    assert (bb.exceptions.NumericalWarning in kinds) == (int(info.max()) > 0)
    assert (bb.exceptions.BotorchWarning in kinds) == (int(info.min()) < 0)
    with warnings.catch_warnings(record=True) as rec:  # drained: nothing is reported twice
        warnings.simplefilter("always")
        acqf.check_numerics()
    assert not rec


def test_optimize_acqf_improves_and_respects_bounds(cuda):
    p = make_problem("C1")
    model = _model(p, cuda)
    acqf = _acqf(p, model, cuda)
    bounds = torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double().to(cuda)
    cand, val = bb.optimize_acqf(acqf, bounds, q=2, num_restarts=4, raw_samples=64,
                                 options={"seed": 0, "maxiter": 20, "batch_limit": 2, "init_batch_limit": 32})  # fmt: skip
    assert cand.shape == (2, p.d) and val.shape == ()
    assert bool(((cand >= 0) & (cand <= 1)).all())
    X_rnd = bb.draw_sobol_samples(bounds.cpu(), n=64, q=2, seed=0)
    with torch.no_grad():
        y_raw = acqf(X_rnd.to(cuda))
    assert float(val) >= float(y_raw.max()) - 1e-12  # the best raw sample is always among the restarts
    assert float(acqf(cand)) == pytest.approx(float(val), rel=1e-12)


def test_gen_candidates_torch_keeps_the_iterate_on_the_device(cuda):
    """gen.py:301-424 on the fused path: Adam over the restarts, no host round trip; the result improves on the start
    and respects the bounds."""
    p = make_problem("C2", n=50)
    model = _model(p, cuda)
    acqf = _acqf(p, model, cuda)
    ics = candidates(p, 6, q=2).to(cuda)
    with torch.no_grad():
        start = acqf(ics)
    cands, vals = bb.gen_candidates_torch(ics, acqf, lower_bounds=0.0, upper_bounds=1.0, options={"maxiter": 40})
    assert cands.is_cuda and cands.shape == ics.shape and not cands.requires_grad
    assert float(cands.min()) >= 0.0 and float(cands.max()) <= 1.0
    assert float(vals.sum()) > float(start.sum())
    best, val = bb.optimize_acqf(acqf, torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double().to(cuda), q=2,
                                 num_restarts=4, raw_samples=32, gen_candidates=bb.gen_candidates_torch,
                                 options={"maxiter": 15, "seed": 0})  # fmt: skip
    assert best.shape == (2, p.d) and bool(torch.isfinite(val))


def test_sequential_greedy_optimize_acqf(cuda):
    """optimize.py:202-243: q greedy single-point optimisations conditioned through X_pending; the i-th reported value is
    the acquisition value of candidate i given the i-1 earlier ones, and X_pending is restored afterwards."""
    p = make_problem("C2", n=50)
    oacq = oracle_acq(p, oracle_gp(p))
    model = _model(p, cuda)
    acqf = _acqf(p, model, cuda)
    bounds = torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double().to(cuda)
    torch.manual_seed(0)  # initialize_q_batch selects restarts with torch.multinomial on the global RNG (initializers.py:946)
    cands, vals = bb.optimize_acqf(acqf, bounds, q=3, num_restarts=4, raw_samples=64, sequential=True,
                                   options={"maxiter": 20, "seed": 0})  # fmt: skip
    assert cands.shape == (3, p.d) and vals.shape == (3,) and acqf.X_pending is None
    for i in range(3):
        ref = oacq.forward(cands[: i + 1].cpu().unsqueeze(0)[:, [i] + list(range(i))])  # new point first, pending after
        # The points come out of an optimiser: greedy candidates sit where the joint covariance with the pending points and
        # the data is close to singular, and there the ORACLE's own rounding noise is ~1e-8 (one run of this test measured
        # 2e-8 against it).  The 1e-9 parity bar is held at seeded points (test_gpu_qlognei.py); this check is about the
        # conditioning logic of sequential greedy, so it only asks for agreement well above that noise.
        assert rel_err(vals[i], ref[0], scale=0.0) < 1e-6
    with pytest.raises(NotImplementedError):
        bb.optimize_acqf(acqf, bounds, q=2, num_restarts=2, raw_samples=8, sequential=True, return_best_only=False)


def test_raw_sample_argmax_index_matches_oracle(cuda):
    p = make_problem("C2")
    oacq = oracle_acq(p, oracle_gp(p))
    acqf = _acqf(p, _model(p, cuda), cuda)
    bounds = torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double()
    X_rnd = bb.draw_sobol_samples(bounds, n=256, q=p.q, seed=5)
    from botorch_b200.optim import evaluate_raw_samples

    y = evaluate_raw_samples(acqf, X_rnd, cuda, batch_limit=100)
    with torch.no_grad():
        ref = torch.cat([oacq.forward(X_rnd[i : i + 64]) for i in range(0, 256, 64)])
    assert int(torch.argmax(y)) == int(torch.argmax(ref))


def test_thompson_sampling_marginal(cuda):
    p = make_problem("C5", n=300)
    ogp = oracle_gp(p)
    model = _model(p, cuda)
    Xc = qmc.sobol_box(torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double(), 2000, 1, seed=1).squeeze(1)
    ts = bb.MaxPosteriorSampling(model, replacement=False, chunk_size=512)
    mean, var = ts.marginal_posterior(Xc.to(cuda))
    m_ref, c_ref = ogp.posterior_direct(Xc.unsqueeze(1))
    assert rel_err(mean[0], m_ref.squeeze(-1)) < RTOL
    assert rel_err(var[0], c_ref.reshape(-1), scale=float(ogp.y_std) ** 2) < RTOL
    g = torch.Generator(device=cuda).manual_seed(0)
    picks = ts(Xc.to(cuda), num_samples=4, generator=g, joint=False)
    assert picks.shape == (4, p.d)
    assert len({tuple(r.tolist()) for r in picks.cpu()}) == 4  # replacement=False
    # a draw with zero noise is the posterior-mean maximiser
    from botorch_b200 import _lib

    val = torch.empty(1, dtype=torch.float64, device=cuda)
    idx = torch.empty(1, dtype=torch.int64, device=cuda)
    _lib.check(_lib.lib().bo_ts_argmax(_lib.ctx(cuda), 1, 2000, mean.data_ptr(), var.data_ptr(), None,
                                       val.data_ptr(), idx.data_ptr(), _lib.stream_ptr(cuda)))  # fmt: skip
    assert int(idx) == int(torch.argmax(m_ref.squeeze(-1)))


def test_thompson_sampling_joint(cuda):
    """generation/sampling.py:95-155: joint posterior draws over the whole candidate set (N x N Cholesky on the GPU),
    then the reference's maximize_samples index logic with and without replacement."""
    p = make_problem("C5", n=200)
    ogp = oracle_gp(p)
    model = _model(p, cuda)
    N = 700
    Xc = qmc.sobol_box(torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double(), N, 1, seed=2).squeeze(1).to(cuda)
    ts = bb.MaxPosteriorSampling(model, replacement=True)
    torch.manual_seed(11)
    picks = ts(Xc, num_samples=5)
    torch.manual_seed(11)
    samples = model.posterior(Xc).rsample(sample_shape=torch.Size([5]))  # 5 x N x 1, same base draws
    assert picks.shape == (5, p.d)
    assert torch.equal(picks, Xc[samples.squeeze(-1).argmax(dim=-1)])
    # the draws are joint: their empirical covariance between two nearby candidates follows the posterior covariance
    torch.manual_seed(3)
    big = model.posterior(Xc[:2]).rsample(sample_shape=torch.Size([4096])).squeeze(-1)
    _, c_ref = ogp.posterior_direct(Xc[:2].cpu())
    emp = torch.cov(big.T).cpu()
    assert float((emp - c_ref).abs().max()) < 0.15 * float(c_ref.abs().max())
    ts_nr = bb.MaxPosteriorSampling(model, replacement=False)
    picks = ts_nr(Xc, num_samples=6)
    assert picks.shape == (6, p.d) and len({tuple(r.tolist()) for r in picks.cpu()}) == 6
    from botorch_b200.generation import _flip_sub_unique

    x = torch.tensor([1, 6, 4, 3, 6, 3])  # generation/utils.py:62-66
    assert _flip_sub_unique(x, 3).tolist() == [3, 6, 4] and _flip_sub_unique(x, 10).tolist() == [3, 6, 4, 1]


def test_sharded_thompson_sampling_is_independent_of_the_sharding(cuda):
    """SURVEY 8e "C5: shard N": per-shard marginal draws + (value, global index) reduction give the picks of the
    unsharded run, for any number of shards, because the base samples are keyed by global candidate position."""
    from botorch_b200 import distributed as bd

    p = make_problem("C5", n=150)
    model = _model(p, cuda)
    N = 40000  # spans three base-sample blocks
    Xc = qmc.sobol_box(torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double(), N, 1, seed=4).squeeze(1).to(cuda)
    for replacement in (True, False):
        ts = bb.MaxPosteriorSampling(model, replacement=replacement, chunk_size=8192)
        full = bd.thompson_sample_sharded(ts, Xc, num_samples=5, seed=7)
        assert full.shape == (5, p.d)
        if not replacement:
            assert len({tuple(r.tolist()) for r in full.cpu()}) == 5
        for ws in (2, 3):
            scorers = [bd._ShardScorer(ts, Xc, *bd.shard_range(N, r, ws)) for r in range(ws)]
            chosen = []
            for s_ in range(5):
                picks = [sc.pick(s_, 7) for sc in scorers]
                g = int(bd._winner(torch.stack([v for v, _ in picks]), torch.stack([i for _, i in picks])))
                chosen.append(g)
                if not replacement:
                    for sc in scorers:
                        sc.exclude(g)
            assert torch.equal(Xc[torch.tensor(chosen, device=cuda)], full)
    z = bd.ts_base_samples(100, 40000, 2, 7, cuda)
    assert torch.equal(z[16284:16294], bd.ts_base_samples(16384, 16394, 2, 7, cuda))  # block-keyed noise


def test_saas_posterior_mixture(cuda):
    p = make_problem("C4", M=4)
    ogp = oracle_gp(p)
    model = _model(p, cuda)
    X = candidates(p, 3, q=2)
    post = model.posterior(X.to(cuda))
    assert post.mean.shape == (3, 4, 2, 1)
    m_ref, c_ref = ogp.posterior_direct(X)  # B x M x q
    assert rel_err(post.mean.squeeze(-1), m_ref) < RTOL
    assert rel_err(post.mixture_mean.squeeze(-1), m_ref.mean(dim=-2)) < RTOL
    assert post.mixture_variance.shape == (3, 2, 1)
    qv = post.quantile(torch.tensor(0.5, device=cuda, dtype=torch.float64))
    assert qv.shape == (3, 2, 1)
    sd = model.state_dict()
    m2 = bb.SaasFullyBayesianSingleTaskGP(p.train_X.to(cuda), p.train_Y.to(cuda))
    m2.load_state_dict(sd)
    assert rel_err(m2.posterior(X.to(cuda)).mean, post.mean) < 1e-9


def test_cuda_path_matches_committed_golden_fixtures(cuda):
    """tests/golden/oracle_configs.pt (minted by tests/golden/make_golden.py): the CUDA path against committed numbers,
    independent of the oracle code that travels with the repo."""
    import os

    gold = torch.load(os.path.join(os.path.dirname(__file__), "golden", "oracle_configs.pt"))
    for tag in ("C1", "C2", "C3", "C4"):
        c = gold[tag]
        p = make_problem(c["cfg"], **c["kw"])
        model = _model(p, cuda)
        acqf = _acqf(p, model, cuda)
        Xd = c["X"].to(cuda).requires_grad_(True)
        val = acqf(Xd)
        (g,) = torch.autograd.grad(val.sum(), Xd)
        assert rel_err(val, c["acq"], scale=0.0) < RTOL, tag
        scale = c["grad"].flatten(1).abs().max(dim=1).values
        assert float(((g.cpu() - c["grad"]).flatten(1).abs().max(dim=1).values / scale).max()) < RTOL, tag
        post = model.posterior(c["X"].to(cuda))
        mean = post.mean.squeeze(-1)
        assert rel_err(mean, c["mean"]) < RTOL, tag
        raw = candidates(p, 128, q=c["q"], seed=c["raw_seed"])
        with torch.no_grad():
            y = acqf(raw.to(cuda))
        assert int(torch.argmax(y)) == c["raw_argmax"], tag
        assert float(y.max()) == pytest.approx(c["raw_max"], rel=1e-9)
    c = gold["C5"]
    p = make_problem("C5", **c["kw"])
    ts = bb.MaxPosteriorSampling(_model(p, cuda))
    mean, var = ts.marginal_posterior(c["X"].to(cuda))
    assert rel_err(mean[0], c["mean"]) < RTOL
    assert int(torch.argmax(mean[0])) == c["argmax_mean"]


def test_fixed_observation_noise_model(cuda):
    """SingleTaskGP(train_X, train_Y, train_Yvar) (gp_regression.py:186-195): K + diag(Yvar / stdv^2); posterior,
    qLogNEI value and gradient on the combined-root path (Sigma(X_b, X) = ys^2 D Khat^-1 K_x) against the oracle."""
    from oracle.gp import GPHyper, OracleGP
    from oracle.qlognei import OracleQLogNEI

    p = make_problem("C2", n=90)
    g = torch.Generator().manual_seed(3)
    Yvar = (0.002 + 0.05 * torch.rand(p.train_Y.shape, generator=g, dtype=torch.float64)) * float(p.train_Y.var())
    h = p.hypers
    model = bb.SingleTaskGP(p.train_X.to(cuda), p.train_Y.to(cuda), Yvar.to(cuda), lengthscale=h["lengthscale"],
                            mean_constant=h["mean"])  # fmt: skip
    stdv = p.train_Y.std(dim=-2)
    hyper = GPHyper("rbf", h["lengthscale"], torch.tensor(0.0, dtype=torch.float64), h["mean"],
                    noise_vec=(Yvar / stdv**2).squeeze(-1))  # fmt: skip
    ogp = OracleGP(p.train_X, p.train_Y, hyper, standardize=True)
    X = candidates(p, 6, q=3)
    Xd = X.to(cuda).requires_grad_(True)
    post = model.posterior(Xd)
    m_ref, c_ref = ogp.posterior_direct(X)
    assert rel_err(post.mean.squeeze(-1), m_ref) < RTOL
    assert rel_err(post.distribution.covariance_matrix, c_ref, scale=float(ogp.y_std) ** 2) < RTOL
    # observation_noise=True adds the mean training noise (models/gpytorch.py:393-400)
    noisy = model.posterior(X.to(cuda), observation_noise=True).variance.squeeze(-1)
    expect = torch.diagonal(c_ref, dim1=-2, dim2=-1) + float(Yvar.mean())
    assert rel_err(noisy, expect, scale=float(ogp.y_std) ** 2) < 1e-8
    oacq = OracleQLogNEI(ogp, p.train_X, num_samples=p.mc_samples, seed=1234)
    acqf = _acqf(p, model, cuda)
    assert acqf._nei.is_train  # the shared-baseline (combined-root) path
    val = acqf(Xd)
    (grad,) = torch.autograd.grad(val.sum(), Xd)
    a_ref, g_ref = oacq.value_and_grad(X)
    assert rel_err(val, a_ref, scale=0.0) < RTOL
    assert rel_err(grad, g_ref) < 1e-8
    with pytest.raises(ValueError):
        bb.SingleTaskGP(p.train_X.to(cuda), p.train_Y.to(cuda), Yvar.to(cuda), noise=0.1)


def test_host_evaluator_direct_route_equals_the_autograd_route(cuda, monkeypatch):
    """optim.HostEvaluator (f_np_wrapper, gen.py:194-222): the direct route (forward + reverse kernels with
    d loss / d acq = -1, preallocated buffers) returns the gradient bit for bit and the loss to rounding of the sum."""
    import numpy as np

    import botorch_b200 as bb
    from botorch_b200.optim import HostEvaluator
    from botorch_b200.synthetic import make_problem
    from tests.util import candidates

    p = make_problem("C2")
    h = p.hypers
    model = bb.SingleTaskGP(p.train_X.to(cuda), p.train_Y.to(cuda), lengthscale=h["lengthscale"], noise=h["noise"],
                            mean_constant=h["mean"])  # fmt: skip
    like = torch.empty(0, dtype=torch.float64, device=cuda)
    for cls, args in ((bb.qLogNoisyExpectedImprovement, (p.train_X.to(cuda),)), (bb.qLogExpectedImprovement, (float(p.train_Y.max()),))):
        acqf = cls(model, *args, sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=3))
        X = candidates(p, 16)
        x = X.numpy().reshape(-1).copy()
        ev = HostEvaluator(acqf, X.shape, like)
        loss_d, grad_d = ev(x)
        loss_d2, grad_d2 = ev(x + 0.0)
        assert ev.direct_calls == 2 and loss_d == loss_d2 and np.array_equal(grad_d, grad_d2)
        monkeypatch.setenv("BOTORCH_B200_DIRECT_EVAL", "0")
        ev_a = HostEvaluator(acqf, X.shape, like)
        loss_a, grad_a = ev_a(x)
        monkeypatch.delenv("BOTORCH_B200_DIRECT_EVAL")
        assert ev_a.direct_calls == 0
        assert np.array_equal(grad_d, grad_a)
        assert abs(loss_d - loss_a) <= 1e-13 * abs(loss_a)
    # SAAS ensemble: the mixture (logmeanexp over the models, utils/transforms.py:289-293) and its softmax weights are formed
    # on the device between the forward and the reverse kernels
    pe = make_problem("C4", n=40, M=5)
    me = bb.SaasFullyBayesianSingleTaskGP(pe.train_X.to(cuda), pe.train_Y.to(cuda))
    me.load_mcmc_samples(pe.hypers)
    for cls, args in ((bb.qLogNoisyExpectedImprovement, (pe.train_X.to(cuda),)), (bb.qNoisyExpectedImprovement, (pe.train_X.to(cuda),))):
        import warnings

        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ae = cls(me, *args, sampler=bb.SobolQMCNormalSampler(torch.Size([64]), seed=3), prune_baseline=False)
        Xe = candidates(pe, 6)
        xe = Xe.numpy().reshape(-1).copy()
        ev_e = HostEvaluator(ae, Xe.shape, like)
        loss_d, grad_d = ev_e(xe)
        assert ev_e.direct_calls == 1
        monkeypatch.setenv("BOTORCH_B200_DIRECT_EVAL", "0")
        loss_a, grad_a = HostEvaluator(ae, Xe.shape, like)(xe)
        monkeypatch.delenv("BOTORCH_B200_DIRECT_EVAL")
        assert abs(loss_d - loss_a) <= 1e-12 * abs(loss_a)
        assert float(np.abs(grad_d - grad_a).max()) <= 1e-12 * float(np.abs(grad_a).max())
    # pending points and float32 callers stay on the general route
    acqf.set_X_pending(candidates(p, 1)[0].to(cuda))
    ev_p = HostEvaluator(acqf, X.shape, like)
    ev_p(x)
    assert ev_p.direct_calls == 0
