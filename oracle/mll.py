"""Oracle restatement of the exact marginal log likelihood (TEST INFRASTRUCTURE, float64 torch CPU).

What ``fit_gpytorch_mll`` evaluates per L-BFGS-B step:

  reference call site                                         restated here
  ----------------------------------------------------------  ---------------------------------------
  fit.py:117-259 ``_fit_fallback``                             (host logic: botorch_b200/fit.py)
  optim/fit.py:47-110 ``fit_gpytorch_mll_scipy``               (host logic)
  optim/closures/model_closures.py:171-184                     ``exact_mll``: model(*train_inputs) in train mode,
     ``-mll(model_output, model.train_targets)``                 likelihood(K) = K + s2 I (or + diag(Yvar)), log_prob
  models/gp_regression.py:249-254 forward                      ConstantMean + ARD kernel on the train inputs
  models/utils/gpytorch_modules.py:74-127                      ``lognormal_log_prob`` priors: lengthscale
                                                                LogNormal(sqrt2 + log(d)/2, sqrt3), noise LogNormal(-4, 1)
  botorch/__init__.py:44-49                                    exact Cholesky (no CG / Lanczos) up to n = 4096

gpytorch==1.12 arithmetic (not vendored, not installable here; PARITY UNPINNED by reference fixtures):
``ExactMarginalLogLikelihood.forward`` = (``MultivariateNormal.log_prob`` + sum of prior log-probs) / n;
``log_prob`` = -1/2 (r^T Khat^-1 r + logdet Khat + n log 2 pi) through the Cholesky factor.
In train mode the lengthscale requires grad, so ``sq_dist`` does NOT zero the diagonal (x1_eq_x2 shortcut is taken only
for inputs without grad): the diagonal is the rounding residue of the augmented GEMM clamped at 0.
Gradients come from torch autograd, like the reference's closure.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor

LOG_2PI = math.log(2.0 * math.pi)


def _sq_dist_train(x: Tensor) -> Tensor:
    """gpytorch.kernels.kernel.sq_dist(x, x, x1_eq_x2=True) for an input that requires grad."""
    adj = x.mean(-2, keepdim=True)
    x = x - adj
    n1 = x.pow(2).sum(dim=-1, keepdim=True)
    p1 = torch.ones_like(n1)
    a = torch.cat([-2.0 * x, n1, p1], dim=-1)
    b = torch.cat([x, p1, n1], dim=-1)
    return a.matmul(b.transpose(-2, -1)).clamp_min(0)


def train_kernel(X: Tensor, kernel: str, lengthscale: Tensor, outputscale: Optional[Tensor]) -> Tensor:
    """covar_module(train_X) in train mode: RBFKernel / MaternKernel(nu=2.5) (x ScaleKernel)."""
    if kernel == "rbf":
        K = _sq_dist_train(X.div(lengthscale)).div(-2).exp()
    elif kernel == "matern52":
        mean = X.reshape(-1, X.size(-1)).mean(0)[(None,) * (X.dim() - 1)]
        dist = _sq_dist_train((X - mean).div(lengthscale)).clamp_min(1e-30).sqrt()
        K = ((math.sqrt(5) * dist).add(1).add(5.0 / 3.0 * dist**2)) * torch.exp(-math.sqrt(5) * dist)
    else:
        raise ValueError(kernel)
    return K if outputscale is None else outputscale * K


def log_marginal(X: Tensor, targets: Tensor, kernel: str, lengthscale: Tensor, noise: Optional[Tensor], mean: Tensor,
                 outputscale: Optional[Tensor] = None, noise_vec: Optional[Tensor] = None) -> Tensor:  # fmt: skip
    """MultivariateNormal(mean, K + D).log_prob(targets), D = s2 I or diag(noise_vec)."""
    n = X.shape[-2]
    K = train_kernel(X, kernel, lengthscale, outputscale)
    D = noise_vec if noise_vec is not None else noise.expand(n)
    L = torch.linalg.cholesky(K + torch.diag_embed(D))
    r = (targets - mean).unsqueeze(-1)
    inv_quad = torch.cholesky_solve(r, L).mul(r).sum()
    logdet = 2.0 * L.diagonal(dim1=-2, dim2=-1).log().sum()
    return -0.5 * (inv_quad + logdet + n * LOG_2PI)


def lognormal_log_prob(x: Tensor, loc: float, scale: float) -> Tensor:
    """torch.distributions.LogNormal(loc, scale).log_prob (gpytorch LogNormalPrior)."""
    lx = x.log()
    return -lx - math.log(scale) - 0.5 * LOG_2PI - (lx - loc) ** 2 / (2.0 * scale * scale)


def gamma_log_prob(x: Tensor, concentration: float, rate: float) -> Tensor:
    """torch.distributions.Gamma(concentration, rate).log_prob (gpytorch GammaPrior)."""
    return concentration * math.log(rate) + (concentration - 1.0) * x.log() - rate * x - math.lgamma(concentration)


def exact_mll(X: Tensor, targets: Tensor, kernel: str, lengthscale: Tensor, noise: Optional[Tensor], mean: Tensor,
              outputscale: Optional[Tensor] = None, noise_vec: Optional[Tensor] = None,
              outputscale_prior: Tuple[float, float] = (2.0, 0.15)) -> Tensor:  # fmt: skip
    """ExactMarginalLogLikelihood of the default SingleTaskGP module set: (log_prob + priors) / n."""
    d, n = X.shape[-1], X.shape[-2]
    res = log_marginal(X, targets, kernel, lengthscale, noise, mean, outputscale, noise_vec)
    res = res + lognormal_log_prob(lengthscale, math.sqrt(2.0) + 0.5 * math.log(d), math.sqrt(3.0)).sum()
    if noise_vec is None:
        res = res + lognormal_log_prob(noise, -4.0, 1.0).sum()
    if outputscale is not None:
        res = res + gamma_log_prob(outputscale, *outputscale_prior).sum()
    return res / n


def value_and_grad(fn, X: Tensor, targets: Tensor, kernel: str, hypers: Dict[str, Optional[Tensor]],
                   noise_vec: Optional[Tensor] = None) -> Tuple[Tensor, Dict[str, Tensor]]:  # fmt: skip
    """fn in (log_marginal, exact_mll); hypers: lengthscale [d], noise [], mean [], outputscale [] or None."""
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in hypers.items() if v is not None}
    if noise_vec is not None:
        leaves.pop("noise", None)
    val = fn(X, targets, kernel, leaves["lengthscale"], leaves.get("noise"), leaves["mean"], leaves.get("outputscale"),
             noise_vec)  # fmt: skip
    grads = torch.autograd.grad(val, list(leaves.values()))
    return val.detach(), dict(zip(leaves.keys(), grads))
