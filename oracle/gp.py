"""Oracle restatement of the exact-GP posterior (TEST INFRASTRUCTURE, float64 torch CPU).

The arithmetic of this part of the path is NOT in /root/reference: it lives in
gpytorch==1.12 and linear_operator==0.5.2 (pinned by requirements.txt:6-7,
neither vendored nor installable here).  This file restates their published
algorithms as they are reached from the reference's call sites:

  reference call site                                   restated here
  ----------------------------------------------------  -------------------------------
  models/gpytorch.py:446  ``mvn = self(X)``              ``OracleGP.posterior``  (ExactGP.__call__ in
                                                          eval mode -> DefaultPredictionStrategy)
  models/gp_regression.py:249-254 forward                ``OracleGP._mean`` / ``_kernel``
  models/utils/gpytorch_modules.py:100-127               RBF-ARD without outputscale (``rbf``)
  models/fully_bayesian.py:81-91, 275-282                ScaleKernel(Matern-5/2 ARD) (``matern52``)
  models/utils/assorted.py:287-297 (fast_pred_var ON)    covariance via the dense ``L^-T`` cache
  botorch/__init__.py:44-49                              exact Cholesky everywhere, 6 jitter tries
  models/transforms/outcome.py:289-307, 420-441          ``Standardize`` fit / ``untransform``
  utils/low_rank.py:141, posteriors/gpytorch.py:121      ``psd_safe_cholesky`` / ``rsample``

PARITY UNPINNED for the values this file produces: the reference's tests hold no
golden vectors for GP posteriors.  ``posterior_direct`` is the
independent second formulation used to bound the oracle's own rounding noise.
"""
from __future__ import annotations

import math
import warnings
from dataclasses import dataclass
from typing import Optional, Tuple

import torch
from torch import Tensor


class NanError(RuntimeError):
    """linear_operator.utils.errors.NanError stand-in."""


class NotPSDError(RuntimeError):
    """linear_operator.utils.errors.NotPSDError stand-in."""


class NumericalWarning(RuntimeWarning):
    """gpytorch.utils.warnings.NumericalWarning stand-in."""


CHOLESKY_MAX_TRIES = 6  # botorch/__init__.py:47
CHOLESKY_JITTER_F64 = 1e-8  # linear_operator settings.cholesky_jitter (double)


def psd_safe_cholesky(A: Tensor, max_tries: int = CHOLESKY_MAX_TRIES, jitter: float = CHOLESKY_JITTER_F64) -> Tensor:
    """linear_operator.utils.cholesky.psd_safe_cholesky (0.5.2).

    Plain potrf first; for batch members that fail, add jitter*10**i to the
    diagonal (i = 0..max_tries-1, cumulative jitter == jitter*10**i) and retry.
    NaN input -> NanError, still failing -> NotPSDError.
    """
    L, info = torch.linalg.cholesky_ex(A)
    if not torch.any(info):
        return L
    if torch.isnan(A).any():
        raise NanError(f"cholesky_cpu: {int(torch.isnan(A).sum())} of {A.numel()} elements are NaN.")
    Ap = A.clone()
    prev = 0.0
    for i in range(max_tries):
        new = jitter * (10**i)
        add = ((info > 0) * (new - prev)).unsqueeze(-1).expand(*Ap.shape[:-1])
        Ap.diagonal(dim1=-1, dim2=-2).add_(add)
        prev = new
        warnings.warn(f"A not p.d., added jitter of {new:.1e} to the diagonal", NumericalWarning)
        L, info = torch.linalg.cholesky_ex(Ap)
        if not torch.any(info):
            return L
    raise NotPSDError(f"Matrix not positive definite after repeatedly adding jitter up to {new:.1e}.")


def sq_dist(x1: Tensor, x2: Tensor, x1_eq_x2: bool = False) -> Tensor:
    """gpytorch.kernels.kernel.sq_dist: centre by x1's mean, augmented GEMM, clamp at 0."""
    adj = x1.mean(-2, keepdim=True)
    x1 = x1 - adj
    n1 = x1.pow(2).sum(dim=-1, keepdim=True)
    p1 = torch.ones_like(n1)
    if x1_eq_x2:
        x2, n2, p2 = x1, n1, p1
    else:
        x2 = x2 - adj
        n2 = x2.pow(2).sum(dim=-1, keepdim=True)
        p2 = torch.ones_like(n2)
    a = torch.cat([-2.0 * x1, n1, p1], dim=-1)
    b = torch.cat([x2, p2, n2], dim=-1)
    res = a.matmul(b.transpose(-2, -1))
    if x1_eq_x2:
        res.diagonal(dim1=-2, dim2=-1).fill_(0)
    return res.clamp_min_(0)


def rbf(x1: Tensor, x2: Tensor, lengthscale: Tensor, x1_eq_x2: bool = False) -> Tensor:
    """gpytorch RBFKernel.forward: exp(-sq_dist(x1/l, x2/l)/2)."""
    return sq_dist(x1.div(lengthscale), x2.div(lengthscale), x1_eq_x2).div(-2).exp()


def matern52(x1: Tensor, x2: Tensor, lengthscale: Tensor, x1_eq_x2: bool = False) -> Tensor:
    """gpytorch MaternKernel.forward (nu=2.5); formula mirrored at models/fully_bayesian.py:81-85."""
    mean = x1.reshape(-1, x1.size(-1)).mean(0)[(None,) * (x1.dim() - 1)]
    a = (x1 - mean).div(lengthscale)
    b = (x2 - mean).div(lengthscale)
    dist = sq_dist(a, b, x1_eq_x2).clamp_min_(1e-30).sqrt_()
    exp_component = torch.exp(-math.sqrt(5) * dist)
    poly = (math.sqrt(5) * dist).add(1).add(5.0 / 3.0 * dist**2)
    return poly * exp_component


@dataclass
class GPHyper:
    """Fitted hyperparameters.  Unbatched (SingleTaskGP) or with a leading M (SAAS MCMC) dim.

    kernel: "rbf" (SingleTaskGP default, no outputscale) or "matern52" (SAAS, with outputscale).
    lengthscale: (d,) or (M, d);  outputscale/noise/mean: scalars () or (M,).
    """

    kernel: str
    lengthscale: Tensor
    noise: Tensor
    mean: Tensor
    outputscale: Optional[Tensor] = None
    # fixed per-observation noise (n,) in the model's (standardised) space: FixedNoiseGaussianLikelihood,
    # /root/reference/botorch/models/gp_regression.py:186-195; the train covariance is then K + diag(noise_vec)
    noise_vec: Optional[Tensor] = None

    @property
    def batch_shape(self) -> torch.Size:
        return self.lengthscale.shape[:-1]


class OracleGP:
    """Exact GP with fixed hyperparameters; posterior as gpytorch's DefaultPredictionStrategy gives it."""

    def __init__(self, train_X: Tensor, train_Y: Tensor, hyper: GPHyper, standardize: bool = True):
        assert train_X.dtype == torch.float64 and train_X.dim() == 2
        self.train_X = train_X
        self.n, self.d = train_X.shape
        self.hyper = hyper
        y = train_Y.reshape(self.n, 1)
        if standardize:  # outcome.py:289-307
            stdv = y.std(dim=-2, keepdim=True) if self.n > 1 else torch.ones(1, 1, dtype=y.dtype)
            stdv = stdv.where(stdv >= 1e-8, torch.full_like(stdv, 1.0))
            self.y_mean, self.y_std = y.mean(dim=-2, keepdim=True).reshape(()), stdv.reshape(())
        else:
            self.y_mean, self.y_std = torch.zeros((), dtype=y.dtype), torch.ones((), dtype=y.dtype)
        self.train_targets = ((y - self.y_mean) / self.y_std).squeeze(-1)  # standardised, (n,)
        self._caches = None

    # -- modules ---------------------------------------------------------------------------
    def _ls(self) -> Tensor:
        return self.hyper.lengthscale.unsqueeze(-2)  # (..., 1, d)

    def _kernel(self, x1: Tensor, x2: Tensor, x1_eq_x2: bool = False) -> Tensor:
        h = self.hyper
        base = (rbf if h.kernel == "rbf" else matern52)(x1, x2, self._ls(), x1_eq_x2)
        if h.outputscale is not None:  # ScaleKernel
            base = base.mul(h.outputscale[..., None, None])
        return base

    def _noise(self) -> Tensor:
        return self.hyper.noise[..., None, None]

    # -- prediction caches (gpytorch DefaultPredictionStrategy.mean_cache / covar_cache) ----
    def caches(self) -> Tuple[Tensor, Tensor, Tensor]:
        if self._caches is None:
            K = self._kernel(self.train_X, self.train_X, x1_eq_x2=True)
            if self.hyper.noise_vec is not None:
                Khat = K + torch.diag_embed(self.hyper.noise_vec)
            else:
                Khat = K + self._noise() * torch.eye(self.n, dtype=K.dtype)
            L = psd_safe_cholesky(Khat)
            resid = (self.train_targets - self.hyper.mean[..., None]).unsqueeze(-1)
            mean_cache = torch.cholesky_solve(resid.expand(*L.shape[:-1], 1), L).squeeze(-1)
            eye = torch.eye(self.n, dtype=K.dtype)
            covar_cache = torch.linalg.solve_triangular(L, eye, upper=False).mT  # L^-T
            self._caches = (L, mean_cache, covar_cache)
        return self._caches

    # -- posterior ---------------------------------------------------------------------------
    def _model_X(self, X: Tensor) -> Tensor:
        """SAAS inserts the MCMC dim at -3 (models/fully_bayesian.py:539)."""
        return X.unsqueeze(-3) if len(self.hyper.batch_shape) > 0 else X

    def posterior_standardized(self, X: Tensor) -> Tuple[Tensor, Tensor]:
        """mean (..., rows), covar (..., rows, rows) in the model's (standardised) space."""
        L, mean_cache, covar_cache = self.caches()
        Xm = self._model_X(X)
        joint = torch.cat([self.train_X.expand(*Xm.shape[:-2], self.n, self.d), Xm], dim=-2)
        K_rows = self._kernel(Xm, joint)  # eager test rows x all columns
        test_train, test_test = K_rows[..., : self.n], K_rows[..., self.n :]
        mean = (test_train @ mean_cache.unsqueeze(-1)).squeeze(-1) + self.hyper.mean[..., None]
        root = test_train @ covar_cache
        covar = torch.add(test_test, root @ root.mT, alpha=-1)
        return mean, covar

    def untransform(self, mean: Tensor, covar: Tensor) -> Tuple[Tensor, Tensor]:
        """Standardize.untransform_posterior (outcome.py:420-441): offset+scale*mean, D covar D."""
        scale = self.y_std.expand_as(mean)
        mean_tf = self.y_mean + self.y_std * mean
        covar_tf = scale.unsqueeze(-1) * covar * scale.unsqueeze(-2)
        return mean_tf, covar_tf

    def posterior(self, X: Tensor) -> Tuple[Tensor, Tensor]:
        """Model.posterior(X) -> (mean, covariance) on the original outcome scale."""
        return self.untransform(*self.posterior_standardized(X))

    def posterior_direct(self, X: Tensor) -> Tuple[Tensor, Tensor]:
        """Independent formulation: K** - K*x (K+s2 I)^-1 Kx* through cholesky_solve."""
        L, mean_cache, _ = self.caches()
        Xm = self._model_X(X)
        Ktx = self._kernel(Xm, self.train_X.expand(*Xm.shape[:-2], self.n, self.d))
        Ktt = self._kernel(Xm, Xm)
        mean = (Ktx @ mean_cache.unsqueeze(-1)).squeeze(-1) + self.hyper.mean[..., None]
        sol = torch.cholesky_solve(Ktx.mT, L)
        covar = Ktt - Ktx @ sol
        return self.untransform(mean, covar)

    def posterior_minimal_blocks(self, X: Tensor, X_b: Tensor):
        """Only the q new rows: mean_q, Sigma(X,X_b), Sigma(X,X) via triangular solves.

        Same mathematics as rows [-q:] of ``posterior(cat[X_b, X])``; this is the
        algorithm the CUDA path implements.
        """
        L, mean_cache, _ = self.caches()
        Xm = self._model_X(X)
        tr = self.train_X.expand(*Xm.shape[:-2], self.n, self.d)
        Kxt = self._kernel(Xm, tr)  # (..., q, n)
        Kxb = self._kernel(Xm, X_b.expand(*Xm.shape[:-2], *X_b.shape[-2:]))
        Kxx = self._kernel(Xm, Xm)
        Kbt = self._kernel(X_b, self.train_X)  # (M?, n_b, n)
        V = torch.linalg.solve_triangular(L, Kxt.mT, upper=False)  # (..., n, q)
        Vb = torch.linalg.solve_triangular(L, Kbt.mT, upper=False)  # (M?, n, n_b)
        mean = (Kxt @ mean_cache.unsqueeze(-1)).squeeze(-1) + self.hyper.mean[..., None]
        S_qb = Kxb - V.mT @ Vb
        S_qq = Kxx - V.mT @ V
        s = self.y_std
        return self.y_mean + s * mean, s * s * S_qb, s * s * S_qq
