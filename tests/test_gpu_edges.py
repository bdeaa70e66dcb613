"""Edge sizes of the fused path against the oracle: smallest and largest q the kernels accept, every q-templated kernel
variant (4 / 8 / 12 / 16 / 32), the widest input dimension, sample counts that are not multiples of anything, a single
training point, a single t-batch -- values and gradients through the C ABI (engine.NEIState)."""
import math

import pytest
import torch

from botorch_b200.synthetic import Problem, default_lengthscale
from oracle import qmc
from tests.test_gpu_qlognei import _grad_rel
from tests.util import gp_state, oracle_acq, oracle_gp, rel_err

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def _problem(n, d, q, S, kernel="rbf", M=None, seed=0):
    g = torch.Generator().manual_seed(seed)
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    Y = torch.sin(3.0 * X.sum(-1, keepdim=True) / math.sqrt(d)) + 0.05 * torch.randn(n, 1, generator=g, dtype=torch.float64)
    if M is None:
        hyp = {"lengthscale": torch.full((d,), default_lengthscale(d), dtype=torch.float64),
               "noise": torch.tensor(1e-2, dtype=torch.float64), "mean": torch.tensor(0.0, dtype=torch.float64)}  # fmt: skip
    else:
        hyp = {"lengthscale": 0.3 + torch.rand(M, d, generator=g, dtype=torch.float64),
               "noise": 1e-3 + 0.05 * torch.rand(M, generator=g, dtype=torch.float64),
               "mean": 0.1 * torch.randn(M, generator=g, dtype=torch.float64),
               "outputscale": 0.5 + torch.rand(M, generator=g, dtype=torch.float64)}  # fmt: skip
    return Problem(f"edge-n{n}-d{d}-q{q}", X, Y, kernel, hyp, q, S, 4, 16, standardize=n > 1)


EDGES = [
    # n,   d,  q,  S,   B, kernel,     M
    (1, 2, 1, 8, 3, "rbf", None),  # one training point, no standardisation
    (7, 3, 17, 5, 2, "rbf", None),  # q above every register-blocked variant but 32 -> generic gram, QT = 32 kernels
    (40, 4, 32, 33, 2, "rbf", None),  # BO_MAX_Q
    (33, 128, 2, 64, 3, "rbf", None),  # BO_MAX_D
    (300, 5, 12, 100, 4, "rbf", None),  # QT = 12 reverse kernel, multi-tile GEMMs
    (65, 6, 16, 257, 1, "rbf", None),  # QT = 16 forward kernel, single t-batch, odd sample count
    (50, 7, 5, 1000, 2, "matern52", 3),  # ensemble, odd q, many samples
    (129, 3, 9, 31, 5, "matern52", 2),
]


@pytest.mark.parametrize("n,d,q,S,B,kernel,M", EDGES)
def test_edge_sizes_match_oracle(cuda, n, d, q, S, B, kernel, M):
    from botorch_b200.engine import NEIState

    p = _problem(n, d, q, S, kernel, M)
    ogp = oracle_gp(p)
    oacq = oracle_acq(p, ogp)
    st = gp_state(p, ogp, cuda)
    nei = NEIState(st, p.train_X.to(cuda), oacq.Z_b.to(cuda))
    bounds = torch.stack([torch.zeros(d), torch.ones(d)]).double()
    X = qmc.sobol_box(bounds, B, q, seed=77) if q * d <= 21201 else torch.rand(B, q, d, dtype=torch.float64)
    Zq = oacq.Z_q(q).to(cuda).contiguous()
    acq, info, work = nei.forward(X.to(cuda), Zq, oacq.tau_relu, oacq.tau_max, with_grad=True)
    assert int(info.min()) >= 0
    a_ref, g_ref = oacq.value_and_grad(X, mode="faithful")
    if st.M == 1:
        a = acq[0]
        g = nei.backward(X.to(cuda), Zq, oacq.tau_relu, oacq.tau_max, torch.ones_like(acq), work)[0]
    else:
        from oracle.safe_math import logmeanexp

        a = logmeanexp(acq.mT.cpu(), dim=-1)
        w = torch.softmax(acq.mT, dim=-1).mT.contiguous()
        g = nei.backward(X.to(cuda), Zq, oacq.tau_relu, oacq.tau_max, w, work).sum(0)
    assert rel_err(a, a_ref, scale=0.0) < RTOL
    assert _grad_rel(g, g_ref) < 1e-8


def test_size_limits_are_reported_not_crashed(cuda):
    """q > 32 and d > 128 are outside the per-candidate kernels: the C ABI answers BO_ERR_UNSUPPORTED (a Python
    exception with the library's message), it never launches out of bounds."""
    from botorch_b200 import _lib
    from botorch_b200.engine import NEIState

    p = _problem(20, 3, 2, 16)
    ogp = oracle_gp(p)
    oacq = oracle_acq(p, ogp)
    st = gp_state(p, ogp, cuda)
    nei = NEIState(st, p.train_X.to(cuda), oacq.Z_b.to(cuda))
    X = torch.rand(2, 33, 3, dtype=torch.float64, device=cuda)
    Zq = torch.randn(16, 33, dtype=torch.float64, device=cuda)
    with pytest.raises(_lib.BoError, match="unsupported"):
        nei.forward(X, Zq, 1e-6, 1e-2, with_grad=False)
    with pytest.raises(Exception):
        gp_state(_problem(10, 129, 1, 8), oracle_gp(_problem(10, 129, 1, 8)), cuda)


def test_failed_cached_update_falls_back_to_standard_sampling(cuda):
    """acquisition/cached_cholesky.py:139-165 / test/acquisition/test_logei.py:672-676: when the cached update fails
    (NanError / NotPSDError) the reference warns ONCE and samples from the joint posterior over [X_baseline; X] instead,
    keeping the cached incumbents.  The device state is corrupted the way the reference test corrupts `_baseline_L`;
    value and gradient must equal the oracle's fallback."""
    import warnings

    import botorch_b200 as bb
    from botorch_b200.exceptions import BotorchWarning
    from botorch_b200.optim import HostEvaluator
    from botorch_b200.synthetic import make_problem
    from tests.util import candidates, oracle_acq, oracle_gp, rel_err

    p = make_problem("C2", n=90)
    h = p.hypers
    oacq = oracle_acq(p, oracle_gp(p))
    model = bb.SingleTaskGP(p.train_X.to(cuda), p.train_Y.to(cuda), lengthscale=h["lengthscale"], noise=h["noise"],
                            mean_constant=h["mean"])  # fmt: skip
    acqf = bb.qLogNoisyExpectedImprovement(model, p.train_X.to(cuda),
                                           sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234))  # fmt: skip
    X = candidates(p, 6, q=3)
    ref_val, ref_grad = oacq.value_and_grad(X, mode="fallback")
    healthy = acqf(X.to(cuda)).cpu()
    assert rel_err(healthy, ref_val, scale=0.0) < 1e-9  # same numbers as long as nothing fails (delta = 0)
    acqf._nei.fwd_cat[:, -5:, :] = float("nan")  # the combined root: the samples of every t-batch turn NaN
    with warnings.catch_warnings(record=True) as ws:
        warnings.simplefilter("always")
        with torch.no_grad():
            val = acqf(X.to(cuda)).cpu()
    assert sum(issubclass(w.category, BotorchWarning) and "Falling back" in str(w.message) for w in ws) == 1
    assert rel_err(val, ref_val, scale=0.0) < 1e-9
    # with a gradient, through the objective gen_candidates_scipy hands to L-BFGS-B: the evaluation is redone
    ev = HostEvaluator(acqf, X.shape, torch.empty(0, dtype=torch.float64, device=cuda))
    with warnings.catch_warnings(record=True) as ws:
        warnings.simplefilter("always")
        loss, grad = ev(X.numpy().reshape(-1).copy())
    assert sum(issubclass(w.category, BotorchWarning) and "Falling back" in str(w.message) for w in ws) == 1
    assert abs(loss + float(ref_val.sum())) < 1e-9 * abs(float(ref_val.sum()))
    g = torch.from_numpy(grad).reshape(X.shape)
    gerr = ((-g - ref_grad).flatten(1).abs().max(1).values / ref_grad.flatten(1).abs().max(1).values).max()
    assert float(gerr) < 1e-9
    # NaN inputs: psd_safe_cholesky of the fallback raises NanError, as in the reference
    from botorch_b200.exceptions import NanError

    Xn = X.clone()
    Xn[0, 0, 0] = float("nan")
    with pytest.raises(NanError), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with torch.no_grad():
            acqf(Xn.to(cuda))


def test_graph_replay_equals_eager_launches_bit_for_bit(cuda, monkeypatch):
    """A recurring small shape is captured into a CUDA graph at its GRAPH_MIN_USES-th use (engine.NEIState.graphed):
    replays must return the bits the eager launch sequence returns, for values and gradients, and one-shot shapes must
    stay eager."""
    import botorch_b200 as bb
    from botorch_b200 import engine
    from botorch_b200.synthetic import make_problem
    from tests.util import candidates

    monkeypatch.setenv("BOTORCH_B200_GRAPH_AFTER", "3")
    p = make_problem("C1")
    h = p.hypers
    model = bb.SingleTaskGP(p.train_X.to(cuda), p.train_Y.to(cuda), lengthscale=h["lengthscale"], noise=h["noise"],
                            mean_constant=h["mean"])  # fmt: skip
    acqf = bb.qLogNoisyExpectedImprovement(model, p.train_X.to(cuda),
                                           sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234))  # fmt: skip
    X = candidates(p, 4).to(cuda)

    def value_and_grad(Xd):
        Xd = Xd.clone().requires_grad_(True)
        v = acqf(Xd)
        (g,) = torch.autograd.grad(v.sum(), Xd)
        return v.detach().clone(), g.clone()

    v0, g0 = value_and_grad(X)  # uses 1 and 2 of the key: eager
    v1, g1 = value_and_grad(X)
    assert not acqf._nei.__dict__.get("_graphs")
    v2, g2 = value_and_grad(X)  # third use: captured and replayed
    v3, g3 = value_and_grad(X)
    graphs = acqf._nei.__dict__.get("_graphs")
    assert graphs and all(isinstance(ge, engine.GraphedEval) for ge in graphs.values())
    for v, g in ((v1, g1), (v2, g2), (v3, g3)):
        assert torch.equal(v, v0) and torch.equal(g, g0)
    X2 = candidates(p, 4, seed=5).to(cuda)
    va, ga = value_and_grad(X2)  # replay with new inputs
    monkeypatch.setenv("BOTORCH_B200_GRAPHS", "0")
    vb, gb = value_and_grad(X2)  # eager
    assert torch.equal(va, vb) and torch.equal(ga, gb)
    assert engine.GRAPH_MIN_USES >= 32  # default: short-lived shapes are never captured (profiles/r02_graph_cost.json)


@pytest.mark.parametrize("cfg,B,reduction", [("C2", 16, "log"), ("C1", 5, "log"), ("C2", 7, "soft")])
def test_local_derivative_mode_equals_the_recomputing_reverse_pass(cuda, cfg, B, reduction):
    """bo_nei.local_derivs: the forward leaves d qmax_s / d f in the sample rows and the reverse kernel scales them, instead
    of re-evaluating log_fatplus / fatmax per (sample, point).  Same value, same gradient up to the association of one
    product; the state with ``keep_samples()`` is the recomputing route (what outcome constraints use)."""
    from botorch_b200 import engine
    from botorch_b200.synthetic import make_problem
    from tests.util import candidates, gp_state, oracle_acq, oracle_gp

    p = make_problem(cfg)
    ogp = oracle_gp(p)
    oacq = oracle_acq(p, ogp)
    st = gp_state(p, ogp, cuda)
    red = engine.REDUCE_LOG if reduction == "log" else engine.REDUCE_LOG_SOFT
    fast = engine.NEIState(st, p.train_X.to(cuda), oacq.Z_b.to(cuda), reduction=red)
    slow = engine.NEIState(st, p.train_X.to(cuda), oacq.Z_b.to(cuda), reduction=red).keep_samples()
    assert fast.c.local_derivs == 1 and slow.c.local_derivs == 0
    X = candidates(p, B).to(cuda)
    Zq = oacq.Z_q(p.q).to(cuda).contiguous()
    gin = torch.linspace(-1.5, 0.5, B, dtype=torch.float64, device=cuda).reshape(1, B).contiguous()
    out = []
    for nei in (fast, slow):
        acq, info, work = nei.forward(X, Zq, oacq.tau_relu, oacq.tau_max, with_grad=True)
        assert int(info.abs().sum()) == 0
        out.append((acq.clone(), nei.backward(X, Zq, oacq.tau_relu, oacq.tau_max, gin, work).clone()))
    (a1, g1), (a0, g0) = out
    assert float((a1 - a0).abs().max()) <= 1e-14 * float(a0.abs().max())
    assert float((g1 - g0).abs().max()) <= 1e-12 * float(g0.abs().max())
    with pytest.raises(Exception):
        fast.samples_view(work, B, p.q)  # the sample rows do not hold samples in this mode


@pytest.mark.parametrize("cfg,general", [("C2", False), ("C2", True), ("C4", False)])
def test_two_phase_baseline_setup_leaves_the_same_bits(cuda, cfg, general):
    """bo_nei.setup_phase: the factors that do not read the base samples (phase 1) are enqueued while the host still draws
    them; phase 2 adds what reads them.  Every buffer must equal the single-call setup bit for bit -- shared baseline
    (combined root), general baseline (X_baseline != train_X), ensemble."""
    from botorch_b200 import engine
    from botorch_b200.synthetic import make_problem
    from tests.util import gp_state, oracle_acq, oracle_gp

    p = make_problem(cfg, n=60) if cfg == "C4" else make_problem(cfg)
    ogp = oracle_gp(p)
    oacq = oracle_acq(p, ogp)
    st = gp_state(p, ogp, cuda)
    Xb, Zb = p.train_X.to(cuda), oacq.Z_b.to(cuda)
    if general:
        Xb, Zb = Xb[:37].contiguous(), Zb[:, :37].contiguous()
    one = engine.NEIState(st, Xb, Zb)
    calls = []

    def produce():
        calls.append(1)
        return Zb.cpu()

    two = engine.NEIState(st, Xb, produce, S=Zb.shape[0])
    assert calls == [1] and two.c.setup_phase == 0
    for a, b in zip(one.factor_buffers(), two.factor_buffers()):
        assert torch.equal(a, b)
    assert torch.equal(one.Z_b, two.Z_b)
    with pytest.raises(ValueError):
        engine.NEIState(st, Xb, produce)  # S is required with a producer
