"""GPU parity of the SAAS potential energy (the function pyro's NUTS evaluates inside
fit_fully_bayesian_model_nuts) against oracle/saas.py, and the batched NUTS around it."""
import pytest
import torch

from oracle import saas as osaas
from tests.util import rel_err

pytestmark = pytest.mark.gpu

TOL = 1e-9


def _problem(n, d, seed):
    g = torch.Generator().manual_seed(seed)
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    y = torch.sin(6.0 * X[:, 0]) + 0.5 * X[:, 1] + 0.1 * torch.randn(n, generator=g, dtype=torch.float64)
    y = (y - y.mean()) / y.std()
    return X, y, g


@pytest.mark.parametrize("n,d,C", [(1, 2, 2), (40, 5, 6), (100, 30, 8), (257, 12, 3)])
def test_saas_potential_matches_oracle(cuda, n, d, C):
    from botorch_b200.fully_bayesian import SaasPotential

    X, y, g = _problem(max(n, 2), d, seed=n)
    X, y = X[:n], y[:n]
    U = 4.0 * torch.rand(C, d + 4, generator=g, dtype=torch.float64) - 2.0
    pot = SaasPotential(X.to(cuda), y.to(cuda).unsqueeze(-1), num_chains=C)
    e, gr = pot(U.to(cuda))
    oe, og = osaas.potential_and_grad(U, X, y)
    assert rel_err(e, oe, scale=float(oe.abs().max())) < TOL
    assert rel_err(gr, og, scale=float(og.abs().max())) < TOL
    c, oc = pot.constrain(U.to(cuda)), osaas.constrained(U, d)
    for k in oc:
        assert rel_err(c[k], oc[k], scale=1.0) < 1e-14


def test_a_divergent_chain_does_not_touch_the_others(cuda):
    from botorch_b200.fully_bayesian import SaasPotential

    X, y, g = _problem(30, 4, seed=2)
    U = 4.0 * torch.rand(5, 8, generator=g, dtype=torch.float64) - 2.0
    pot = SaasPotential(X.to(cuda), y.to(cuda), num_chains=5)
    ref_e, ref_g = pot(U.to(cuda))
    V = U.clone()
    V[1] = float("nan")
    V[3, 0] = 500.0  # exp overflows
    e, gr = pot(V.to(cuda))
    assert bool(torch.isinf(e[[1, 3]]).all()) and bool((gr[[1, 3]] == 0).all())
    keep = [0, 2, 4]
    assert torch.equal(e[keep], ref_e[keep]) and torch.equal(gr[keep], ref_g[keep])
    oe, _ = osaas.potential_and_grad(U[keep], X, y)
    assert rel_err(e[keep], oe, scale=float(oe.abs().max())) < TOL


def test_graph_replay_equals_the_eager_evaluation(cuda):
    """From the third call on the potential is one CUDA-graph replay over static buffers: same numbers, bit for bit, also
    when a chain of the batch diverges."""
    from botorch_b200.fully_bayesian import SaasPotential

    X, y, g = _problem(100, 6, seed=3)
    C = 4
    eager = SaasPotential(X.to(cuda), y.to(cuda), num_chains=C, graph=False)
    graphed = SaasPotential(X.to(cuda), y.to(cuda), num_chains=C)
    for i in range(6):
        U = (4.0 * torch.rand(C, 10, generator=g, dtype=torch.float64) - 2.0).to(cuda)
        if i == 4:
            U[2] = float("nan")
        e0, g0 = eager(U)
        e1, g1 = graphed(U)
        assert torch.equal(e0, e1) and torch.equal(g0, g1)
        if i == 4:
            assert bool(torch.isinf(e1[2]))
    assert graphed._graph is not None and eager._graph is None


def test_nuts_on_the_device_potential_follows_the_oracle_potential(cuda):
    """Same sampler, same device RNG stream; the energy comes from bo_mll_value_grad in one run and from the CPU oracle in the
    other: the first draws must coincide (differences only grow through the chaotic dynamics)."""
    from botorch_b200.fully_bayesian import SaasPotential, nuts_sample

    X, y, g = _problem(25, 3, seed=5)
    C = 3
    init = (torch.rand(C, 7, generator=g, dtype=torch.float64) - 0.5).to(cuda)
    pot = SaasPotential(X.to(cuda), y.to(cuda), num_chains=C)

    def oracle_pot(u):
        e, gr = osaas.potential_and_grad(u.cpu(), X, y, divergent_as_inf=True)
        return e.to(cuda), gr.to(cuda)

    a, sa = nuts_sample(pot, init, num_samples=4, warmup_steps=3, max_tree_depth=4, seed=11, adapt_mass=False)
    b, sb = nuts_sample(oracle_pot, init, num_samples=4, warmup_steps=3, max_tree_depth=4, seed=11, adapt_mass=False)
    assert sa["leapfrogs"] == sb["leapfrogs"]
    assert rel_err(a, b, scale=1.0) < 1e-6
    assert not torch.equal(a[0], a[-1])


def test_graphed_subtrees_draw_what_the_eager_sampler_draws(cuda):
    """Each subtree of 2**depth leapfrog steps as one CUDA-graph launch: bit-identical draws and statistics."""
    from botorch_b200.fully_bayesian import SaasPotential, nuts_sample

    X, y, g = _problem(40, 5, seed=8)
    C = 4
    init = (torch.rand(C, 9, generator=g, dtype=torch.float64) - 0.5).to(cuda)
    runs = []
    for graph in (False, True):
        pot = SaasPotential(X.to(cuda), y.to(cuda), num_chains=C)
        runs.append(nuts_sample(pot, init, num_samples=6, warmup_steps=45, max_tree_depth=4, seed=2, graph=graph) + (pot,))
    (a, sa, pa), (b, sb, pb) = runs
    assert torch.equal(a, b) and sa["leapfrogs"] == sb["leapfrogs"] and pa.evaluations == pb.evaluations
    assert torch.equal(sa["step_size"], sb["step_size"]) and torch.equal(sa["inverse_mass"], sb["inverse_mass"])


def test_fit_fully_bayesian_model_nuts_end_to_end(cuda):
    import botorch_b200 as bb

    X, y, _ = _problem(40, 4, seed=9)  # only the first two inputs matter
    model = bb.SaasFullyBayesianSingleTaskGP(X.to(cuda), y.unsqueeze(-1).to(cuda))
    with pytest.raises(RuntimeError):
        model.posterior(X[:3].to(cuda))
    bb.fit_fully_bayesian_model_nuts(model, max_tree_depth=5, warmup_steps=96, num_samples=64, thinning=4, num_chains=4, seed=0)
    assert model.num_mcmc_samples == 16
    ls = model.median_lengthscale
    assert ls.shape == (4,) and bool(torch.isfinite(ls).all())
    assert float(ls[0]) < float(ls[2:].min())  # the sparsity prior switches the irrelevant inputs off
    post = model.posterior(X[:5].to(cuda))
    assert post.mean.shape == (16, 5, 1) and bool(torch.isfinite(post.mean).all())
    mixture = post.mean.mean(0).reshape(-1)
    assert float((mixture.cpu() - y[:5]).abs().max()) < 0.6
    st = model.nuts_stats
    assert 0.5 < float(st["accept_prob"].mean()) <= 1.0 and st["potential_evaluations"] > 100
    acqf = bb.qLogNoisyExpectedImprovement(model, X.to(cuda), sampler=bb.SobolQMCNormalSampler(torch.Size([16]), seed=0))
    val = acqf(torch.rand(3, 2, 4, dtype=torch.float64, device=cuda))
    assert val.shape == (3,) and bool(torch.isfinite(val).all())
