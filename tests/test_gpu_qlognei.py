"""GPU parity of the fused qLogNEI forward/backward (through the C ABI) against the CPU oracle.

Tolerances (BASELINE.json north_star): acquisition values and gradients within 1e-9 relative in
float64; raw-sample argmax index bit-exact.  "Relative" for the gradient is relative to the largest
component of that t-batch element's gradient (components that are ~0 carry no relative meaning).
"""
import pytest
import torch

from botorch_b200.synthetic import make_problem
from oracle import qmc
from tests.util import candidates, gp_state, oracle_acq, oracle_gp, rel_err

pytestmark = pytest.mark.gpu

RTOL = 1e-9


def _setup(cuda, cfg, kw, X_baseline=None):
    from botorch_b200.engine import NEIState

    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    oacq = oracle_acq(p, ogp, X_baseline=X_baseline)
    st = gp_state(p, ogp, cuda)
    Xb = p.train_X if X_baseline is None else X_baseline
    nei = NEIState(st, Xb.to(cuda), oacq.Z_b.to(cuda))
    return p, ogp, oacq, st, nei


def _grad_rel(g, g_ref):
    g, g_ref = g.detach().cpu(), g_ref.detach().cpu()
    scale = g_ref.flatten(1).abs().max(dim=1).values.clamp_min(1e-300)
    return float(((g - g_ref).flatten(1).abs().max(dim=1).values / scale).max())


CASES = [
    ("C1", {}, 1, 16),
    ("C1", {}, 3, 7),
    ("C2", {}, 4, 32),
    ("C2", {"n": 131}, 2, 5),
    ("C3", {"n": 300}, 8, 12),
    ("C4", {"M": 6}, 4, 9),
    # the inferred-noise floor 1e-4 (gpytorch_modules.py:29): SURVEY.md 8(d)'s ill-conditioned stress case
    ("C1", {"noise": 1e-4}, 2, 6),
    ("C2", {"noise": 1e-4}, 4, 8),
    ("C2", {"n": 400, "noise": 1e-4}, 3, 6),
    ("C3", {"n": 300, "noise": 1e-4}, 8, 6),
]


@pytest.mark.parametrize("cfg,kw,q,B", CASES)
def test_baseline_state_matches_oracle(cuda, cfg, kw, q, B):
    p, ogp, oacq, st, nei = _setup(cuda, cfg, kw)
    Lb = oacq.baseline_L
    assert rel_err(nei.baseline_L.reshape(Lb.shape) if Lb.dim() == 2 else nei.baseline_L, Lb) < 1e-9
    bf = oacq.best_f  # S x (M)
    bf = bf.unsqueeze(-1) if bf.dim() == 1 else bf
    assert rel_err(nei.best_f.mT, bf) < 1e-9


@pytest.mark.parametrize("cfg,kw,q,B", CASES)
def test_forward_backward_match_oracle(cuda, cfg, kw, q, B):
    p, ogp, oacq, st, nei = _setup(cuda, cfg, kw)
    X = candidates(p, B, q=q)
    Zq = oacq.Z_q(q)
    Xd = X.to(cuda)
    acq, info, work = nei.forward(Xd, Zq.to(cuda).contiguous(), oacq.tau_relu, oacq.tau_max, with_grad=True)
    assert int(info.abs().sum()) == 0
    # per-model values: the oracle reduces over the MCMC dim at the very end, so compare after that reduction too
    a_ref, g_ref = oacq.value_and_grad(X, mode="faithful")
    M = st.M
    if M == 1:
        a = acq[0]
        gin = -torch.ones_like(acq)  # loss = -acq.sum(), as gen_candidates_scipy builds it
        g = -nei.backward(Xd, Zq.to(cuda).contiguous(), oacq.tau_relu, oacq.tau_max, gin, work)[0]
    else:
        from oracle.safe_math import logmeanexp

        a = logmeanexp(acq.mT.cpu(), dim=-1)
        w = torch.softmax(acq.mT, dim=-1).mT.contiguous()  # d logmeanexp / d acq[m, b]
        g = nei.backward(Xd, Zq.to(cuda).contiguous(), oacq.tau_relu, oacq.tau_max, w, work).sum(0)
    assert rel_err(a, a_ref, scale=0.0) < RTOL
    assert _grad_rel(g, g_ref) < RTOL


def test_forward_no_grad_equals_with_grad_and_is_deterministic(cuda):
    p, ogp, oacq, st, nei = _setup(cuda, "C2", {})
    X = candidates(p, 40).to(cuda)
    Zq = oacq.Z_q(p.q).to(cuda).contiguous()
    a1, _, _ = nei.forward(X, Zq, 1e-6, 1e-2, with_grad=False)
    a2, _, _ = nei.forward(X, Zq, 1e-6, 1e-2, with_grad=True)
    a3, _, _ = nei.forward(X, Zq, 1e-6, 1e-2, with_grad=False)
    assert torch.equal(a1, a2) and torch.equal(a1, a3)
    # a candidate's value does not depend on its position in the t-batch (shared base samples)
    perm = torch.randperm(40, generator=torch.Generator().manual_seed(0)).to(cuda)
    a4, _, _ = nei.forward(X[perm].contiguous(), Zq, 1e-6, 1e-2, with_grad=False)
    assert torch.equal(a4, a1[:, perm])


@pytest.mark.parametrize("cfg,kw,q,B,kind", [
    ("C2", {}, 4, 9, "subset"), ("C3", {"n": 300}, 8, 5, "subset"), ("C2", {"n": 90}, 3, 6, "off-train"),
    ("C4", {"M": 4}, 4, 6, "subset"), ("C4", {"M": 3}, 2, 5, "off-train"),
])  # fmt: skip
def test_separate_baseline_set(cuda, cfg, kw, q, B, kind):
    """X_baseline != train_X (a pruned baseline, or points that were never observed): extra cross-covariance
    K(X_b, X) and its gradient path, RBF and Matern, single- and multi-tile sizes."""
    p = make_problem(cfg, **kw)
    Xb = p.train_X[::3].contiguous()
    if kind == "off-train":
        Xb = torch.cat([Xb[:7], candidates(p, 6, q=1, seed=21).reshape(-1, p.d)])
    p, ogp, oacq, st, nei = _setup(cuda, cfg, kw, X_baseline=Xb)
    assert not nei.is_train
    X = candidates(p, B, q=q)
    Zq = oacq.Z_q(q).to(cuda).contiguous()
    acq, info, work = nei.forward(X.to(cuda), Zq, 1e-6, 1e-2, with_grad=True)
    a_ref, g_ref = oacq.value_and_grad(X, mode="faithful")
    if st.M == 1:
        a = acq[0]
        g = nei.backward(X.to(cuda), Zq, 1e-6, 1e-2, torch.ones_like(acq), work)[0]
    else:
        from oracle.safe_math import logmeanexp

        a = logmeanexp(acq.mT.cpu(), dim=-1)
        w = torch.softmax(acq.mT, dim=-1).mT.contiguous()
        g = nei.backward(X.to(cuda), Zq, 1e-6, 1e-2, w, work).sum(0)
    assert rel_err(a, a_ref, scale=0.0) < RTOL
    assert _grad_rel(g, g_ref) < RTOL


def test_raw_sample_argmax_is_bit_exact(cuda):
    """optim/initializers.py:944: the index of the best raw sample must equal the oracle's."""
    from botorch_b200 import _lib

    p, ogp, oacq, st, nei = _setup(cuda, "C2", {})
    X = candidates(p, 512)
    Zq = oacq.Z_q(p.q).to(cuda).contiguous()
    acq, _, _ = nei.forward(X.to(cuda), Zq, 1e-6, 1e-2, with_grad=False)
    with torch.no_grad():
        ref = torch.cat([oacq.forward(X[i : i + 64], mode="faithful") for i in range(0, 512, 64)])
    val = torch.empty(1, dtype=torch.float64, device=cuda)
    idx = torch.empty(1, dtype=torch.int64, device=cuda)
    _lib.check(_lib.lib().bo_argmax(_lib.ctx(cuda), 512, acq.data_ptr(), val.data_ptr(), idx.data_ptr(),
                                    _lib.stream_ptr(cuda)))  # fmt: skip
    assert int(idx) == int(torch.argmax(ref)) == int(torch.argmax(acq[0]))
    assert float(val) == float(acq[0].max())


def test_candidate_on_training_point_takes_jitter_path(cuda):
    """A candidate equal to a baseline point makes the Schur block singular: psd_safe_cholesky's jitter
    schedule must kick in (info > 0 => NumericalWarning on the host), values stay finite."""
    p, ogp, oacq, st, nei = _setup(cuda, "C1", {})
    X = p.train_X[:4].reshape(4, 1, 2).contiguous()
    Zq = oacq.Z_q(1).to(cuda).contiguous()
    acq, info, _ = nei.forward(X.to(cuda), Zq, 1e-6, 1e-2, with_grad=False)
    assert bool(torch.isfinite(acq).all())
    assert int(info.min()) >= 0
