"""Posterior objects returned by the fused models.

Same contract as the reference's posterior classes, without the lazy linear_operator graph:
  * ``Posterior``                 botorch/posteriors/posterior.py
  * ``FusedGPPosterior``          botorch/posteriors/gpytorch.py:35-179 (``GPyTorchPosterior``, single output)
  * ``GaussianMixturePosterior``  botorch/posteriors/fully_bayesian.py:82-158 (MCMC dim at -3)

The mean vector and the dense q x q covariance block come straight out of the CUDA kernels
(``bo_posterior_joint``); nothing else is materialised.  Sampling factors the block with the blocked
Cholesky of libbotorch_b200 and applies the base samples with its triangular GEMM.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor

from . import engine

MCMC_DIM = -3  # position of the hyperparameter-sample dimension of a fully Bayesian posterior
_BISECT_TOL = 1e-6


class DenseMVN:
    """Just enough of ``gpytorch.distributions.MultivariateNormal`` for the callers on the path."""

    def __init__(self, mean: Tensor, covariance_matrix: Tensor) -> None:
        self.loc = mean
        self.covariance_matrix = covariance_matrix
        self._root: Optional[Tensor] = None

    mean = property(lambda self: self.loc)
    lazy_covariance_matrix = property(lambda self: self.covariance_matrix)
    batch_shape = property(lambda self: self.loc.shape[:-1])
    event_shape = property(lambda self: self.loc.shape[-1:])
    base_sample_shape = property(lambda self: self.loc.shape[-1:])

    @property
    def variance(self) -> Tensor:
        return self.covariance_matrix.diagonal(dim1=-2, dim2=-1)

    @property
    def stddev(self) -> Tensor:
        return self.variance.sqrt()

    def root(self) -> Tensor:
        """Cholesky root of the covariance, cached like LinearOperator.root_decomposition."""
        if self._root is None:
            self._root = engine.mvn_root(self.covariance_matrix.detach())
        return self._root

    def rsample(self, sample_shape: torch.Size = torch.Size(), base_samples: Optional[Tensor] = None) -> Tensor:
        sample_shape = torch.Size(sample_shape)
        full = sample_shape + self.loc.shape
        if base_samples is None:
            base_samples = torch.randn(full, dtype=self.loc.dtype, device=self.loc.device)
        flat = base_samples.expand(full).reshape(max(sample_shape.numel(), 1), *self.loc.shape)
        if torch.is_grad_enabled() and (self.loc.requires_grad or self.covariance_matrix.requires_grad):
            out = engine.root_sample(self.loc, self.covariance_matrix, flat)  # differentiable w.r.t. X through mean / cov
        else:
            out = engine.root_apply(self.root(), self.loc.detach(), flat)
        return out.reshape(full).to(self.loc.dtype)


class Posterior:
    """Abstract posterior (names as in botorch/posteriors/posterior.py)."""

    def rsample(self, sample_shape: Optional[torch.Size] = None) -> Tensor:  # pragma: no cover
        raise NotImplementedError

    def sample(self, sample_shape: Optional[torch.Size] = None) -> Tensor:
        with torch.no_grad():
            return self.rsample(sample_shape=sample_shape)

    @property
    def batch_range(self) -> Tuple[int, int]:  # pragma: no cover
        raise NotImplementedError


class FusedGPPosterior(Posterior):
    """Single-output Gaussian posterior over q points per t-batch."""

    _is_mt = False

    def __init__(self, distribution: DenseMVN) -> None:
        self.distribution = distribution

    # `posterior.mvn` is what older call sites read (acquisition/cached_cholesky.py:119)
    mvn = property(lambda self: self.distribution)

    def __getattr__(self, name: str):  # attribute fall-through (posteriors/torch.py:74-79)
        dist = self.__dict__.get("distribution")
        if dist is None or name.startswith("__"):
            raise AttributeError(name)
        return getattr(dist, name)

    @property
    def device(self) -> torch.device:
        return self.distribution.loc.device

    @property
    def dtype(self) -> torch.dtype:
        return self.distribution.loc.dtype

    @property
    def base_sample_shape(self) -> torch.Size:
        return self.distribution.batch_shape + self.distribution.base_sample_shape

    @property
    def batch_range(self) -> Tuple[int, int]:
        return (0, -1)

    @property
    def batch_shape(self) -> torch.Size:
        return self.distribution.batch_shape

    def _extended_shape(self, sample_shape: torch.Size = torch.Size()) -> torch.Size:
        return torch.Size(sample_shape) + self.distribution.batch_shape + self.distribution.event_shape + torch.Size([1])

    @property
    def mean(self) -> Tensor:
        return self.distribution.mean.unsqueeze(-1)

    @property
    def variance(self) -> Tensor:
        return self.distribution.variance.unsqueeze(-1)

    def rsample_from_base_samples(self, sample_shape: torch.Size, base_samples: Tensor) -> Tensor:
        if base_samples.shape[: len(sample_shape)] != sample_shape:
            raise RuntimeError(
                "`sample_shape` disagrees with shape of `base_samples`. "
                f"Got {sample_shape=} and {base_samples.shape=}."
            )
        return self.distribution.rsample(sample_shape=sample_shape, base_samples=base_samples).unsqueeze(-1)

    def rsample(self, sample_shape: Optional[torch.Size] = None) -> Tensor:
        sample_shape = torch.Size([1]) if sample_shape is None else sample_shape
        return self.distribution.rsample(sample_shape=sample_shape).unsqueeze(-1)

    def quantile(self, value: Tensor) -> Tensor:
        if value.numel() > 1:
            return torch.stack([self.quantile(v) for v in value], dim=0)
        return torch.distributions.Normal(self.mean, self.variance.sqrt()).icdf(value)

    def density(self, value: Tensor) -> Tensor:
        if value.numel() > 1:
            return torch.stack([self.density(v) for v in value], dim=0)
        return torch.distributions.Normal(self.mean, self.variance.sqrt()).log_prob(value).exp()


def _bisect(f, target: float, bounds: Tensor, tol: float = _BISECT_TOL, max_steps: int = 32) -> Tensor:
    lo_ok, hi_ok = f(bounds[0]) <= target, target <= f(bounds[1])
    if not bool((lo_ok & hi_ok).all()):
        raise ValueError("The target is not contained in the interval specified by the bounds")
    lo, hi = bounds[0].clone(), bounds[1].clone()
    mid = 0.5 * (lo + hi)
    for _ in range(max_steps):
        fm = f(mid)
        if float((fm - target).abs().max()) <= tol:
            break
        above = fm > target
        hi = torch.where(above, mid, hi)
        lo = torch.where(above, lo, mid)
        mid = 0.5 * (lo + hi)
    return mid


class GaussianMixturePosterior(FusedGPPosterior):
    """Equal-weight mixture over the hyperparameter samples at dimension ``MCMC_DIM``."""

    def __init__(self, distribution: DenseMVN) -> None:
        super().__init__(distribution)
        self._mean = distribution.mean.unsqueeze(-1)
        self._variance = distribution.variance.unsqueeze(-1)
        self._covariance_matrix = distribution.covariance_matrix
        self._mix_mean = self._mix_var = self._mix_cov = None

    @property
    def mixture_mean(self) -> Tensor:
        if self._mix_mean is None:
            self._mix_mean = self._mean.mean(dim=MCMC_DIM)
        return self._mix_mean

    @property
    def mixture_variance(self) -> Tensor:
        if self._mix_var is None:  # law of total variance over the mixture components
            m = self._mean.shape[MCMC_DIM]
            within = self._variance.sum(dim=MCMC_DIM) / m
            second = self._mean.pow(2).sum(dim=MCMC_DIM) / m
            self._mix_var = within + second - (self._mean.sum(dim=MCMC_DIM) / m).pow(2)
        return self._mix_var

    @property
    def mixture_covariance_matrix(self) -> Tensor:
        if self._mix_cov is None:
            m = self._mean.shape[MCMC_DIM]
            dev = self._mean - self.mixture_mean.unsqueeze(MCMC_DIM)
            self._mix_cov = (self._covariance_matrix.sum(dim=MCMC_DIM) + (dev @ dev.mT).sum(dim=MCMC_DIM)) / m
        return self._mix_cov

    def quantile(self, value: Tensor) -> Tensor:
        if value.numel() > 1:
            return torch.stack([self.quantile(v) for v in value], dim=0)
        if value <= 0 or value >= 1:
            raise ValueError("value is expected to be in the range (0, 1).")
        comp = torch.distributions.Normal(loc=self.mean, scale=self.variance.sqrt())
        icdf = comp.icdf(value)
        if self.mean.shape[MCMC_DIM] == 1:
            return icdf.squeeze(MCMC_DIM)
        bounds = torch.stack([icdf.min(dim=MCMC_DIM).values - _BISECT_TOL, icdf.max(dim=MCMC_DIM).values + _BISECT_TOL])
        return _bisect(lambda x: comp.cdf(x.unsqueeze(MCMC_DIM)).mean(dim=MCMC_DIM), float(value), bounds)
