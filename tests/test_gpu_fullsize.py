"""GPU parity at the FULL sizes of BASELINE.json's configs C3, C4 and C5 (through the C ABI, against the CPU oracle).

The small-n cases of test_gpu_qlognei.py exercise 5-row-tile problems; the benchmarked shapes are a different code
regime (n = 2000: 36 x 64 row tiles, `tri_off` = 264, LPT tile order, 4 CTAs/SM; M = 256 batched models; n = 20 000
blocked factorisation with 313 diagonal blocks), so they are tested at size here.  Tolerances are north_star's: 1e-9
relative for values, gradients and means; covariances / variances 1e-9 of the prior variance; argmax indices
bit-exact.  Each assertion message carries the oracle's own floor (its two formulations against each other).
"""
import time

import pytest
import torch

from botorch_b200.synthetic import make_problem
from tests.util import candidates, gp_state, oracle_acq, oracle_gp, rel_err

pytestmark = pytest.mark.gpu

RTOL = 1e-9


def _grad_rel(g, g_ref):
    g, g_ref = g.detach().cpu(), g_ref.detach().cpu()
    scale = g_ref.flatten(1).abs().max(dim=1).values.clamp_min(1e-300)
    return float(((g - g_ref).flatten(1).abs().max(dim=1).values / scale).max())


def _oracle_chunks(oacq, X, mode, chunk):
    vals, grads = [], []
    for i in range(0, X.shape[0], chunk):
        v, g = oacq.value_and_grad(X[i : i + chunk], mode=mode)
        vals.append(v)
        grads.append(g)
    return torch.cat(vals), torch.cat(grads)


def test_c3_full_size_forward_backward(cuda):
    """BASELINE.json configs[2]: Ackley-20, n = 2000, q = 8, S = 256; B = 64 t-batches (8 column tiles of the stacked
    GEMMs, all 36 row tiles).  Oracle: minimal algorithm on all 64, the reference's order of operations (posterior over
    [X_baseline; X], L^-T GEMM covariance: utils/low_rank.py:118-162) on 4 of them."""
    from botorch_b200.engine import NEIState

    p = make_problem("C3")
    assert (p.n, p.d, p.q, p.mc_samples) == (2000, 20, 8, 256)
    ogp = oracle_gp(p)
    oacq = oracle_acq(p, ogp)
    st = gp_state(p, ogp, cuda)
    nei = NEIState(st, p.train_X.to(cuda), oacq.Z_b.to(cuda))
    B = 64
    X = candidates(p, B)
    Zq = oacq.Z_q(p.q).to(cuda).contiguous()
    Xd = X.to(cuda)
    acq, info, work = nei.forward(Xd, Zq, oacq.tau_relu, oacq.tau_max, with_grad=True)
    assert int(info.abs().sum()) == 0
    g = nei.backward(Xd, Zq, oacq.tau_relu, oacq.tau_max, torch.ones_like(acq), work)[0]
    a_min, g_min = _oracle_chunks(oacq, X, "minimal", 8)
    a_fth, g_fth = _oracle_chunks(oacq, X[:4], "faithful", 2)
    floor_v = rel_err(a_min[:4], a_fth, scale=0.0)
    floor_g = _grad_rel(g_min[:4], g_fth)
    msg = f"oracle floor (minimal vs faithful): value {floor_v:.1e}, gradient {floor_g:.1e}"
    assert rel_err(acq[0], a_min, scale=0.0) < RTOL, msg
    assert _grad_rel(g, g_min) < RTOL, msg
    assert rel_err(acq[0, :4], a_fth, scale=0.0) < RTOL, msg
    assert _grad_rel(g[:4], g_fth) < RTOL, msg
    # baseline state at size
    assert rel_err(nei.baseline_L[0], oacq.baseline_L) < RTOL
    assert rel_err(nei.best_f[0], oacq.best_f) < RTOL
    # bit-exact argmax over the 64 values (optim/initializers.py:944)
    assert int(torch.argmax(acq[0])) == int(torch.argmax(a_min))


def test_c4_full_size_forward_backward(cuda):
    """BASELINE.json configs[3]: SAAS ensemble, M = 256 hyperparameter samples, n = 100, d = 30, q = 4, S = 256; B = 8
    t-batches; mixture reduction logmeanexp over the MCMC dim (utils/transforms.py:289-293).  Oracle: the reference's
    order of operations on the whole ensemble."""
    from botorch_b200.engine import NEIState
    from oracle.safe_math import logmeanexp

    p = make_problem("C4")
    assert (p.n, p.d, p.q, p.mc_samples) == (100, 30, 4, 256) and p.hypers["lengthscale"].shape[0] == 256
    ogp = oracle_gp(p)
    oacq = oracle_acq(p, ogp)
    st = gp_state(p, ogp, cuda)
    assert st.M == 256
    nei = NEIState(st, p.train_X.to(cuda), oacq.Z_b.to(cuda))
    B = 8
    X = candidates(p, B)
    Zq = oacq.Z_q(p.q).to(cuda).contiguous()
    Xd = X.to(cuda)
    acq, info, work = nei.forward(Xd, Zq, oacq.tau_relu, oacq.tau_max, with_grad=True)
    assert int((info < 0).sum()) == 0
    a = logmeanexp(acq.mT.cpu(), dim=-1)
    w = torch.softmax(acq.mT, dim=-1).mT.contiguous()
    g = nei.backward(Xd, Zq, oacq.tau_relu, oacq.tau_max, w, work).sum(0)
    a_fth, g_fth = _oracle_chunks(oacq, X, "faithful", 4)
    a_min, g_min = _oracle_chunks(oacq, X, "minimal", 4)
    msg = (f"oracle floor (minimal vs faithful): value {rel_err(a_min, a_fth, scale=0.0):.1e}, "
           f"gradient {_grad_rel(g_min, g_fth):.1e}")  # fmt: skip
    assert rel_err(a, a_fth, scale=0.0) < RTOL, msg
    assert _grad_rel(g, g_fth) < RTOL, msg


def test_c5_full_size_posterior_and_thompson_argmax(cuda):
    """BASELINE.json configs[4]: exact GP n = 20 000, d = 12, marginal Thompson posterior (generation/sampling.py:95-117)
    on a 2048-candidate subsample of the 1M-candidate stream (SURVEY 8c golden-vector list).  Oracle: K + s2 I factored on
    the host (one 20 000^3 / 3 Cholesky) and `K** - K*x (K + s2 I)^-1 Kx*` through cholesky_solve
    (OracleGP.posterior_direct's formulation), built from oracle/gp.py's kernel.  Mean 1e-9 relative, variance 1e-9 of the
    prior variance, Thompson argmax index bit-exact."""
    import botorch_b200 as bb
    from oracle.gp import rbf

    p = make_problem("C5")
    assert (p.n, p.d) == (20000, 12)
    Nc = 2048
    Xc = torch.quasirandom.SobolEngine(p.d, scramble=True, seed=1).draw(Nc, dtype=torch.float64)
    h = p.hypers
    t0 = time.perf_counter()
    model = bb.SingleTaskGP(p.train_X.to(cuda), p.train_Y.to(cuda), lengthscale=h["lengthscale"], noise=h["noise"],
                            mean_constant=h["mean"])  # fmt: skip
    ts = bb.MaxPosteriorSampling(model, chunk_size=Nc)
    mean, var = ts.marginal_posterior(Xc.to(cuda))
    torch.cuda.synchronize()
    t_gpu = time.perf_counter() - t0
    assert float(model.gp_state.jitter.abs().max()) == 0.0
    # ---- host oracle (float64 torch CPU; memory ~10 GB) ----
    t0 = time.perf_counter()
    y = p.train_Y.squeeze(-1)
    y_mean, y_std = y.mean(), y.std()
    targets = (y - y_mean) / y_std
    ls = h["lengthscale"].unsqueeze(-2)
    K = rbf(p.train_X, p.train_X, ls, x1_eq_x2=True)
    K.diagonal().add_(h["noise"])
    L = torch.linalg.cholesky(K)
    del K
    alpha = torch.cholesky_solve((targets - h["mean"]).unsqueeze(-1), L).squeeze(-1)
    Ktx = rbf(Xc, p.train_X, ls)  # Nc x n
    m_ref = y_mean + y_std * (Ktx @ alpha + h["mean"])
    V = torch.linalg.solve_triangular(L, Ktx.mT, upper=False)  # n x Nc
    v_ref_tri = y_std**2 * (1.0 - (V * V).sum(0))
    sol = torch.cholesky_solve(Ktx.mT, L)
    v_ref_dir = y_std**2 * (1.0 - (Ktx * sol.mT).sum(-1))
    del sol, V, L
    t_cpu = time.perf_counter() - t0
    prior = float(y_std**2)
    floor = float((v_ref_tri - v_ref_dir).abs().max()) / prior
    msg = f"oracle floor (triangular vs cholesky_solve variance): {floor:.1e} of the prior; gpu {t_gpu:.1f} s, host {t_cpu:.1f} s"
    assert rel_err(mean[0], m_ref) < RTOL, msg
    assert rel_err(var[0], v_ref_dir, scale=prior) < RTOL, msg
    assert rel_err(var[0], v_ref_tri, scale=prior) < RTOL, msg
    # Thompson draw + maximize_samples (sampling.py:115-117) with shared base samples: index bit-exact
    from botorch_b200 import _lib

    g = torch.Generator().manual_seed(5)
    for _ in range(3):
        z = torch.randn(Nc, generator=g, dtype=torch.float64)
        ref_idx = int(torch.argmax(m_ref + v_ref_dir.clamp_min(0).sqrt() * z))
        val = torch.empty(1, dtype=torch.float64, device=cuda)
        idx = torch.empty(1, dtype=torch.int64, device=cuda)
        zd = z.to(cuda)
        _lib.check(_lib.lib().bo_ts_argmax(_lib.ctx(cuda), 1, Nc, mean.data_ptr(), var.data_ptr(), zd.data_ptr(),
                                           val.data_ptr(), idx.data_ptr(), _lib.stream_ptr(cuda)))  # fmt: skip
        assert int(idx) == ref_idx, msg
    # the sharded pick equals the unsharded pick at this size (8e "C5: shard N"): base samples keyed by global position
    from botorch_b200 import distributed as bd

    Xd = Xc.to(cuda)
    one = bd.thompson_sample_sharded(ts, Xd, num_samples=2, seed=3, shard=(0, 1))
    parts = [bd._ShardScorer(ts, Xd, *bd.shard_range(Nc, r, 4)) for r in range(4)]
    for s in range(2):
        picks = [sc.pick(s, 3) for sc in parts]
        V_, I_ = torch.stack([p_[0] for p_ in picks]), torch.stack([p_[1] for p_ in picks])
        assert torch.equal(Xd[bd._winner(V_, I_)].reshape(-1), one[s].reshape(-1))
