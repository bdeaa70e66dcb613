"""Oracle restatement of pathwise (Matheron) posterior sampling (TEST INFRASTRUCTURE, float64 torch CPU).

  reference                                                        restated here
  ---------------------------------------------------------------  -----------------------------------------
  sampling/pathwise/features/generators.py:66-110, 113-155           ``fourier_features``: sqrt(os) sqrt(2/F) [sin, cos](x/l . W^T)
  sampling/pathwise/utils.py:84-117 (SineCosine / InverseLengthscale /
     Outputscale transforms)
  sampling/pathwise/prior_samplers.py:51-91                          ``prior_paths``: m + phi(x) . w
  sampling/pathwise/update_strategies.py:72-103                      ``update_weights``: cholesky_solve(y - f0(X) - sqrt(D) eps)
  sampling/pathwise/posterior_samplers.py:143-199, paths.py          ``matheron_paths``: prior + k(x, X) v, Standardize.untransform
  acquisition/thompson_sampling.py:60-87                             evaluating the paths on a candidate set

The covariance function k and the Cholesky factor are gpytorch arithmetic (oracle/gp.py; PARITY UNPINNED by fixtures).
The random draws (W, w, eps) are inputs, produced by the reference's own generators (draw_sobol_normal_samples, randn).
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor

from .gp import matern52, rbf


def fourier_features(X: Tensor, W: Tensor, lengthscale: Tensor, outputscale: Optional[Tensor]) -> Tensor:
    """[N x F] features: scale * cat([sin(z), cos(z)]), z = (X / l) W^T."""
    z = (lengthscale.reciprocal() * X) @ W.transpose(-1, -2)
    F = 2 * W.shape[0]
    phi = (2.0 / F) ** 0.5 * torch.cat([z.sin(), z.cos()], dim=-1)
    return phi if outputscale is None else outputscale.sqrt() * phi


def matheron_paths(X: Tensor, train_X: Tensor, targets: Tensor, kernel: str, lengthscale: Tensor, noise: Tensor,
                   mean: Tensor, outputscale: Optional[Tensor], W: Tensor, w: Tensor, eps: Tensor, y_mean: float = 0.0,
                   y_std: float = 1.0, noise_vec: Optional[Tensor] = None) -> Tensor:  # fmt: skip
    """Path values [S x N] at X [N x d] on the outcome scale."""
    kfun = rbf if kernel == "rbf" else matern52
    osc = 1.0 if outputscale is None else outputscale
    n = train_X.shape[0]
    D = noise_vec if noise_vec is not None else noise.expand(n)
    K = osc * kfun(train_X, train_X, lengthscale, x1_eq_x2=True) + torch.diag_embed(D)
    L = torch.linalg.cholesky(K)
    f0_train = mean + w @ fourier_features(train_X, W, lengthscale, outputscale).T  # [S x n]
    sample_values = f0_train + D.sqrt() * eps
    v = torch.cholesky_solve((targets - sample_values).unsqueeze(-1), L).squeeze(-1)  # [S x n]
    f0 = mean + w @ fourier_features(X, W, lengthscale, outputscale).T
    upd = v @ (osc * kfun(train_X, X, lengthscale))
    return y_mean + y_std * (f0 + upd)
