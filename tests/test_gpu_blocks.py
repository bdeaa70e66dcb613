"""GPU parity of the building blocks exported by the C ABI (DMMA GEMM, blocked Cholesky / inverse,
covariance matrix, GP setup) against torch fp64 / the oracle.  All calls go through libbotorch_b200."""
import ctypes as C

import pytest
import torch

from tests.util import candidates, gp_state, oracle_gp, rel_err

pytestmark = pytest.mark.gpu


def _dgemm(dev, A, B, C0=None, alpha=1.0, beta=0.0, tri=0):
    from botorch_b200 import _lib

    batch, M, K = A.shape
    N = B.shape[-1]
    lda, ldb, ldc = _lib.pad8(K), _lib.pad8(N), _lib.pad8(N)
    Ap = torch.zeros(batch, M, lda, dtype=torch.float64, device=dev)
    Ap[:, :, :K] = A
    Bp = torch.zeros(batch, K, ldb, dtype=torch.float64, device=dev)
    Bp[:, :, :N] = B
    Cp = torch.zeros(batch, M, ldc, dtype=torch.float64, device=dev)
    if C0 is not None:
        Cp[:, :, :N] = C0
    _lib.check(
        _lib.lib().bo_dgemm(_lib.ctx(dev), M, N, K, batch, alpha, Ap.data_ptr(), lda, M * lda, Bp.data_ptr(), ldb,
                            K * ldb, beta, Cp.data_ptr(), ldc, M * ldc, tri, _lib.stream_ptr(dev))
    )  # fmt: skip
    torch.cuda.synchronize()
    return Cp[:, :, :N]


@pytest.mark.parametrize(
    "batch,M,N,K",
    [(1, 8, 8, 4), (1, 1, 1, 1), (3, 37, 29, 53), (1, 130, 257, 100), (2, 64, 64, 64), (1, 513, 1100, 777),
     (1, 2001, 4096, 2000), (5, 101, 256, 100), (1, 256, 1, 300)],
)  # fmt: skip
def test_dgemm_matches_torch(cuda, batch, M, N, K):
    g = torch.Generator(device="cpu").manual_seed(M * 1000 + N)
    A = torch.randn(batch, M, K, generator=g, dtype=torch.float64).to(cuda)
    B = torch.randn(batch, K, N, generator=g, dtype=torch.float64).to(cuda)
    C0 = torch.randn(batch, M, N, generator=g, dtype=torch.float64).to(cuda)
    out = _dgemm(cuda, A, B, C0, alpha=-0.5, beta=2.0)
    ref = -0.5 * (A @ B) + 2.0 * C0
    assert rel_err(out, ref) < 1e-13


@pytest.mark.parametrize("n,N", [(50, 40), (200, 300), (700, 129)])
def test_dgemm_triangular_modes(cuda, n, N):
    g = torch.Generator(device="cpu").manual_seed(n)
    A = torch.randn(1, n, n, generator=g, dtype=torch.float64).to(cuda)
    B = torch.randn(1, n, N, generator=g, dtype=torch.float64).to(cuda)
    lo, up = torch.tril(A), torch.triu(A)
    assert rel_err(_dgemm(cuda, lo, B, tri=1), lo @ B) < 1e-13
    assert rel_err(_dgemm(cuda, up, B, tri=2), up @ B) < 1e-13
    Bl = torch.tril(torch.randn(1, n, n, generator=g, dtype=torch.float64)).to(cuda)
    assert rel_err(_dgemm(cuda, A, Bl, tri=3), A @ Bl) < 1e-13


@pytest.mark.parametrize("n,batch", [(1, 1), (5, 2), (64, 1), (65, 3), (100, 4), (200, 1), (777, 1), (2000, 1)])
def test_potrf_and_inverse(cuda, n, batch):
    from botorch_b200 import _lib

    g = torch.Generator(device="cpu").manual_seed(n)
    R = torch.randn(batch, n, n, generator=g, dtype=torch.float64)
    A = (R @ R.mT / n + 0.5 * torch.eye(n, dtype=torch.float64)).to(cuda)
    lda = _lib.pad8(n)
    Ap = torch.zeros(batch, n, lda, dtype=torch.float64, device=cuda)
    Ap[:, :, :n] = A
    inv = torch.zeros_like(Ap)
    info = torch.zeros(batch, dtype=torch.int32, device=cuda)
    wb = _lib.lib().bo_potrf_workspace_bytes(n, batch)
    work = _lib.workspace(wb, cuda)
    _lib.check(_lib.lib().bo_potrf(_lib.ctx(cuda), n, batch, Ap.data_ptr(), lda, n * lda, inv.data_ptr(),
                                   work.data_ptr(), work.numel(), info.data_ptr(), _lib.stream_ptr(cuda)))  # fmt: skip
    torch.cuda.synchronize()
    assert int(info.abs().sum()) == 0
    L = Ap[:, :, :n]
    Lref = torch.linalg.cholesky(A)
    assert rel_err(L, Lref) < 1e-12
    assert float(torch.triu(L, 1).abs().max()) == 0.0
    eye = torch.eye(n, dtype=torch.float64, device=cuda)
    assert float((inv[:, :, :n] @ Lref - eye).abs().max()) < 1e-11


def test_potrf_reports_first_bad_pivot(cuda):
    from botorch_b200 import _lib

    n = 70
    A = torch.eye(n, dtype=torch.float64, device=cuda).repeat(2, 1, 1)
    A[1, 66, 66] = -1.0
    lda = _lib.pad8(n)
    Ap = torch.zeros(2, n, lda, dtype=torch.float64, device=cuda)
    Ap[:, :, :n] = A
    info = torch.zeros(2, dtype=torch.int32, device=cuda)
    work = _lib.workspace(_lib.lib().bo_potrf_workspace_bytes(n, 2), cuda)
    _lib.check(_lib.lib().bo_potrf(_lib.ctx(cuda), n, 2, Ap.data_ptr(), lda, n * lda, None, work.data_ptr(),
                                   work.numel(), info.data_ptr(), _lib.stream_ptr(cuda)))  # fmt: skip
    assert info.tolist() == [0, 67]  # LAPACK convention: 1-based index of the failing leading minor


@pytest.mark.parametrize("cfg,kw", [("C1", {}), ("C2", {}), ("C4", {"M": 5}), ("C3", {"n": 300})])
def test_kernel_matrix_and_factors(cuda, cfg, kw):
    from botorch_b200.synthetic import make_problem

    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    st = gp_state(p, ogp, cuda)
    Xc = candidates(p, 33).reshape(-1, p.d)
    K = st.kernel_matrix(Xc.to(cuda))  # [M, n, Nc]
    Kref = ogp._kernel(Xc.unsqueeze(-3) if ogp.hyper.batch_shape else Xc, p.train_X).mT
    assert rel_err(K.reshape(Kref.shape) if not ogp.hyper.batch_shape else K, Kref) < 1e-13
    L, mean_cache, covar_cache = ogp.caches()
    assert rel_err(st.L.reshape(L.shape) if L.dim() == 2 else st.L, L) < 1e-11
    assert rel_err(st.alpha.reshape(mean_cache.shape), mean_cache) < 1e-10
    assert rel_err(st.L_inv.reshape(L.shape) if L.dim() == 2 else st.L_inv, covar_cache.mT) < 1e-10


@pytest.mark.parametrize("cfg,kw,q", [("C1", {}, 1), ("C2", {}, 4), ("C4", {"M": 5}, 3), ("C3", {"n": 300}, 8)])
def test_posterior_joint_and_marginal(cuda, cfg, kw, q):
    from botorch_b200.synthetic import make_problem

    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    st = gp_state(p, ogp, cuda)
    X = candidates(p, 17, q=q)
    mean, cov = st.posterior_joint(X.to(cuda))  # [M,B,q], [M,B,q,q]
    m_ref, c_ref = ogp.posterior(X)  # [B,(M),q], [B,(M),q,q]
    if ogp.hyper.batch_shape:
        m_ref, c_ref = m_ref.transpose(0, 1), c_ref.transpose(0, 1)
    prior = float(ogp.y_std**2 * (ogp.hyper.outputscale.max() if ogp.hyper.outputscale is not None else 1.0))
    assert rel_err(mean.reshape(m_ref.shape), m_ref) < 1e-9
    # covariance: relative to the prior variance scale (cancellation K** - |L^-1 k|^2)
    assert rel_err(cov.reshape(c_ref.shape), c_ref, scale=prior) < 1e-9
    mv, vv = st.posterior_mean_var(X.reshape(-1, p.d).to(cuda))
    assert rel_err(mv.reshape(m_ref.shape), m_ref) < 1e-9
    assert rel_err(vv.reshape(m_ref.shape), c_ref.diagonal(dim1=-1, dim2=-2), scale=prior) < 1e-9
