"""GPU parity of pathwise (Matheron) posterior samples (bo_paths_setup / bo_paths_eval) against the
oracle restatement, plus the properties that pin the construction: interpolation of the data as the noise vanishes,
agreement of the path ensemble with the exact posterior, and the Thompson argmax over a candidate set."""
import pytest
import torch

from oracle import pathwise as opw
from tests.util import rel_err

pytestmark = pytest.mark.gpu

TOL = 1e-9


def _problem(n, d, kernel, seed, noise=0.02):
    g = torch.Generator().manual_seed(seed)
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    y = torch.sin(5.0 * X[:, 0]) + X[:, -1] ** 2 + 0.05 * torch.randn(n, generator=g, dtype=torch.float64)
    ls = 0.3 + torch.rand(d, generator=g, dtype=torch.float64)
    h = {"lengthscale": ls, "noise": torch.tensor(noise, dtype=torch.float64), "mean": torch.tensor(0.1, dtype=torch.float64),
         "outputscale": torch.tensor(1.3, dtype=torch.float64) if kernel == "matern52" else None}  # fmt: skip
    return X, y, h


def _draws(S, F, d, n, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(F // 2, d, generator=g, dtype=torch.float64), torch.randn(S, F, generator=g, dtype=torch.float64),
            torch.randn(S, n, generator=g, dtype=torch.float64))  # fmt: skip


@pytest.mark.parametrize("n,d,kernel,S,F,N", [(40, 3, "rbf", 4, 64, 50), (150, 6, "matern52", 3, 256, 1000),
                                             (300, 12, "rbf", 1, 1024, 777), (65, 2, "rbf", 16, 128, 9)])  # fmt: skip
def test_paths_match_oracle(cuda, n, d, kernel, S, F, N):
    from botorch_b200.engine import GPState, PathState

    X, y, h = _problem(n, d, kernel, seed=n)
    ymean, ystd = float(y.mean()), float(y.std())
    t = (y - ymean) / ystd
    W, w, eps = _draws(S, F, d, n, seed=n + 1)
    Xc = torch.rand(N, d, dtype=torch.float64, generator=torch.Generator().manual_seed(3))
    gp = GPState(X.to(cuda), t.to(cuda), kernel, h["lengthscale"].to(cuda), h["noise"].to(cuda), h["mean"].to(cuda),
                 None if h["outputscale"] is None else h["outputscale"].to(cuda), y_mean=ymean, y_std=ystd)  # fmt: skip
    ps = PathState(gp, W.to(cuda), w.to(cuda), eps.to(cuda))
    got = ps.evaluate(Xc.to(cuda)).cpu()
    ref = opw.matheron_paths(Xc, X, t, kernel, h["lengthscale"], h["noise"], h["mean"], h["outputscale"], W, w, eps, ymean, ystd)
    assert got.shape == (S, N)
    assert rel_err(got, ref, scale=float(ref.abs().max())) < TOL
    # Thompson pick per path: same argmax index
    assert torch.equal(got.argmax(dim=-1), ref.argmax(dim=-1))


def test_zero_draws_give_the_posterior_mean(cuda):
    """With w = 0 and eps = 0 a path is m + k(x, X) (K + s2 I)^-1 (y - m): exactly the posterior mean of the model."""
    from botorch_b200 import SingleTaskGP, draw_matheron_paths
    from botorch_b200.engine import PathState

    X, y, h = _problem(60, 2, "rbf", seed=8)
    model = SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["lengthscale"], noise=1e-2)
    st = model.gp_state
    F = 32
    ps = PathState(st, torch.randn(F // 2, 2, dtype=torch.float64).to(cuda), torch.zeros(2, F, dtype=torch.float64).to(cuda),
                   torch.zeros(2, 60, dtype=torch.float64).to(cuda))  # fmt: skip
    Xc = torch.rand(33, 2, dtype=torch.float64, generator=torch.Generator().manual_seed(4)).to(cuda)
    mean = model.posterior(Xc.unsqueeze(0)).mean.reshape(-1)
    vals = ps.evaluate(Xc)
    assert rel_err(vals[0], mean) < TOL and rel_err(vals[1], mean) < TOL
    # the public constructor: sample_shape leads, the input's leading dims follow
    torch.manual_seed(0)
    paths = draw_matheron_paths(model, torch.Size([5]), num_features=512)
    assert paths(X.to(cuda)).shape == (5, 60)
    assert paths(X.to(cuda).reshape(6, 10, 2)).shape == (5, 6, 10)
    with pytest.raises(Exception):
        draw_matheron_paths(model, torch.Size([2]), num_features=33)


def test_path_ensemble_matches_exact_posterior_moments(cuda):
    from botorch_b200 import SingleTaskGP, draw_matheron_paths

    X, y, h = _problem(80, 3, "rbf", seed=5, noise=0.05)
    model = SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["lengthscale"], noise=0.05)
    Xc = torch.rand(40, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(1)).to(cuda)
    torch.manual_seed(1)
    vals = draw_matheron_paths(model, torch.Size([2048]), num_features=2048)(Xc)
    post = model.posterior(Xc.unsqueeze(0))
    mean, var = post.mean.reshape(-1), post.variance.reshape(-1)
    sd = var.sqrt().clamp_min(1e-3)
    assert float(((vals.mean(0) - mean).abs() / sd).max()) < 0.15  # ~ 6 / sqrt(2048) standard errors
    assert float((vals.var(0) / var.clamp_min(1e-6) - 1.0).abs().median()) < 0.15  # RFF prior + Monte-Carlo error


def test_pathwise_thompson_sampling_over_a_candidate_set(cuda):
    from botorch_b200 import MaxPosteriorSampling, SingleTaskGP

    X, y, h = _problem(100, 4, "rbf", seed=12)
    model = SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["lengthscale"], noise=0.01)
    cand = torch.rand(50000, 4, dtype=torch.float64, generator=torch.Generator().manual_seed(2)).to(cuda)
    ts = MaxPosteriorSampling(model, pathwise=True, num_features=256, chunk_size=8192)
    torch.manual_seed(3)
    picks = ts(cand, num_samples=4)
    assert picks.shape == (4, 4)
    # every pick is a row of the candidate set, and good picks beat the average candidate under the posterior mean
    assert all(bool((cand == p).all(dim=-1).any()) for p in picks)
    mean_all = model.posterior(cand[:4096].unsqueeze(0)).mean.mean()
    assert float(model.posterior(picks.unsqueeze(0)).mean.mean()) > float(mean_all)
    ts_nr = MaxPosteriorSampling(model, pathwise=True, num_features=256, replacement=False)
    torch.manual_seed(3)
    picks_nr = ts_nr(cand[:2000], num_samples=3)
    assert len({tuple(p.tolist()) for p in picks_nr}) == 3


def test_sharded_pathwise_thompson_is_sharding_invariant(cuda):
    """Every shard draws the same paths (same host seed); winners over 1 and over 3 shards coincide."""
    from botorch_b200 import MaxPosteriorSampling, SingleTaskGP
    from botorch_b200.distributed import _winner, pathwise_shard_scores, shard_range, thompson_sample_sharded

    X, y, h = _problem(80, 3, "rbf", seed=14)
    model = SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["lengthscale"], noise=0.01)
    cand = torch.rand(10001, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(6)).to(cuda)
    ts = MaxPosteriorSampling(model, pathwise=True, num_features=128, chunk_size=4096)
    single = thompson_sample_sharded(ts, cand, num_samples=5, seed=11, shard=(0, 1))
    parts = [pathwise_shard_scores(ts, cand, 5, 11, *shard_range(cand.shape[0], r, 3)) for r in range(3)]
    win = _winner(torch.stack([p[0] for p in parts]), torch.stack([p[1] for p in parts]))
    assert torch.equal(cand[win], single)
    state = torch.random.get_rng_state()
    thompson_sample_sharded(ts, cand, num_samples=2, seed=3, shard=(0, 1))
    assert torch.equal(state, torch.random.get_rng_state())  # the caller's host RNG stream is left untouched


@pytest.mark.parametrize("n,d,kernel,S,F,N", [(40, 3, "rbf", 4, 64, 50), (150, 6, "matern52", 3, 256, 300),
                                             (600, 5, "rbf", 2, 128, 1100), (65, 1, "rbf", 8, 32, 9)])  # fmt: skip
def test_path_gradient_matches_oracle_autograd(cuda, n, d, kernel, S, F, N):
    from botorch_b200.engine import GPState, PathState, path_values

    X, y, h = _problem(n, d, kernel, seed=n + 3)
    ymean, ystd = float(y.mean()), float(y.std())
    t = (y - ymean) / ystd
    W, w, eps = _draws(S, F, d, n, seed=n + 5)
    g = torch.Generator().manual_seed(9)
    Xc = torch.rand(N, d, dtype=torch.float64, generator=g)
    coef = torch.randn(S, N, dtype=torch.float64, generator=g)
    gp = GPState(X.to(cuda), t.to(cuda), kernel, h["lengthscale"].to(cuda), h["noise"].to(cuda), h["mean"].to(cuda),
                 None if h["outputscale"] is None else h["outputscale"].to(cuda), y_mean=ymean, y_std=ystd)  # fmt: skip
    ps = PathState(gp, W.to(cuda), w.to(cuda), eps.to(cuda))
    Xg = Xc.to(cuda).requires_grad_(True)
    vals = path_values(ps, Xg)
    (grad,) = torch.autograd.grad((vals * coef.to(cuda)).sum(), Xg)
    Xr = Xc.clone().requires_grad_(True)
    ref = opw.matheron_paths(Xr, X, t, kernel, h["lengthscale"], h["noise"], h["mean"], h["outputscale"], W, w, eps, ymean, ystd)
    (gref,) = torch.autograd.grad((ref * coef).sum(), Xr)
    assert rel_err(grad, gref, scale=float(gref.abs().max())) < TOL


def test_pathwise_thompson_sampling_acquisition_under_optimize_acqf(cuda):
    from botorch_b200 import PathwiseThompsonSampling, SingleTaskGP, gen_candidates_torch, optimize_acqf

    X, y, h = _problem(50, 2, "rbf", seed=17)
    model = SingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda), lengthscale=h["lengthscale"], noise=0.01)
    torch.manual_seed(5)
    ts = PathwiseThompsonSampling(model, num_features=256)
    Xq = torch.rand(7, 3, 2, dtype=torch.float64, generator=torch.Generator().manual_seed(1)).to(cuda)
    v = ts(Xq)
    assert v.shape == (7,) and ts.batch_size == 3
    # path j scores point j: the value is the sum of the three paths on their own points
    vals = ts.samples(Xq)  # [3 x 7 x 3]
    assert torch.allclose(v, sum(vals[j, :, j] for j in range(3)), rtol=1e-12, atol=1e-12)
    with pytest.raises(ValueError):
        ts(Xq[:, :2])
    bounds = torch.stack([torch.zeros(2), torch.ones(2)]).double().to(cuda)
    cand, val = optimize_acqf(ts, bounds, q=3, num_restarts=4, raw_samples=64, gen_candidates=gen_candidates_torch,
                              options={"maxiter": 40, "seed": 0})  # fmt: skip
    assert cand.shape == (3, 2) and bool(((cand >= 0) & (cand <= 1)).all())
    raw_best = float(ts(torch.rand(64, 3, 2, dtype=torch.float64, device=cuda)).max())
    assert float(val) >= raw_best - 1e-6 or float(val) > float(v.max())
