"""Oracle restatement of qLogNoisyExpectedImprovement (TEST INFRASTRUCTURE, float64 torch CPU).

Follows
  * /root/reference/botorch/acquisition/logei.py:364-420   _init_baseline (cache_root=True)
  * /root/reference/botorch/acquisition/logei.py:448-481   _get_samples_and_objectives
  * /root/reference/botorch/acquisition/cached_cholesky.py:94-165  root decomposition / _get_f_X_samples
  * /root/reference/botorch/utils/low_rank.py:85-173        sample_cached_cholesky
  * /root/reference/botorch/acquisition/monte_carlo.py:253-289  forward / _non_reduced_forward
  * /root/reference/botorch/utils/transforms.py:289-293     ensemble logmeanexp over the MCMC dim
Gradients come from torch.autograd exactly as in the reference (generation/gen.py:209-211).
"""
from __future__ import annotations

import torch
from torch import Tensor

from . import qmc, safe_math
from .gp import NanError, OracleGP, psd_safe_cholesky

TAU_RELU = 1e-6  # logei.py:66
TAU_MAX = 1e-2  # logei.py:67


class OracleQLogNEI:
    def __init__(
        self,
        model: OracleGP,
        X_baseline: Tensor,
        num_samples: int = 512,
        seed: int = 1234,
        tau_relu: float = TAU_RELU,
        tau_max: float = TAU_MAX,
        obj_weight: float = 1.0,
        fat: bool = True,
    ):
        self.model = model
        self.fat = bool(fat)  # logei.py:122,532: fatmax / log_fatplus, or smooth_amax / log_softplus
        self.obj_weight = float(obj_weight)  # one-outcome LinearMCObjective (objective.py:217-256): obj = w * samples
        self.X_baseline = X_baseline
        self.S = num_samples
        self.seed = seed
        self.tau_relu, self.tau_max = tau_relu, tau_max
        self.n_b = X_baseline.shape[-2]
        self.is_ensemble = len(model.hyper.batch_shape) > 0
        self._Zq_cache = {}
        with torch.no_grad():  # logei.py:391-419
            mean_b, cov_b = model.posterior(X_baseline)  # (M?, n_b), (M?, n_b, n_b)
            self.Z_b = qmc.sobol_normal(d=self.n_b, n=self.S, seed=seed)  # S x n_b
            self.baseline_L = psd_safe_cholesky(cov_b)  # cached root (cached_cholesky.py:115-120)
            # MVN.rsample with base samples: root @ z + loc
            samples = torch.einsum("...ij,sj->s...i", self.baseline_L, self.Z_b) + mean_b
            self.baseline_samples = samples  # S x (M?) x n_b
            self.best_f = (self.obj_weight * samples).max(dim=-1).values  # S x (M?)   (acquisition/utils.py:125-129)

    # -- base samples ----------------------------------------------------------------------
    def Z_q(self, q: int) -> Tensor:
        if q not in self._Zq_cache:
            self._Zq_cache[q] = qmc.qlognei_base_samples(self.S, self.n_b, q, self.seed)[1]
        return self._Zq_cache[q]

    # -- samples at the q new points --------------------------------------------------------
    def _samples_from_blocks(self, mean_q: Tensor, bl: Tensor, br: Tensor) -> Tensor:
        """low_rank.py:118-162 given the bottom rows [bl | br] of the joint covariance."""
        q = br.shape[-1]
        bl_chol = torch.linalg.solve_triangular(self.baseline_L, bl.mT, upper=False).mT
        br_chol = psd_safe_cholesky(br - bl_chol @ bl_chol.mT)
        new_Lq = torch.cat([bl_chol, br_chol], dim=-1)  # (..., q, n_b+q)
        Z = torch.cat([self.Z_b, self.Z_q(q)], dim=-1)  # S x (n_b+q)
        res = torch.einsum("...ij,sj->s...i", new_Lq, Z) + mean_q  # S x ... x q
        if torch.isnan(res).any() or torch.isinf(res).any():
            raise NanError("Samples contain nans or infs.")
        return res

    def samples_faithful(self, X: Tensor) -> Tensor:
        """Reference order of operations: posterior over [X_baseline; X], keep the last q rows."""
        q = X.shape[-2]
        Xb = self.X_baseline.expand(*X.shape[:-2], *self.X_baseline.shape[-2:])
        mean, cov = self.model.posterior(torch.cat([Xb, X], dim=-2))
        bottom = cov[..., -q:, :]
        bl, br = bottom.split([bottom.shape[-1] - q, q], dim=-1)
        return self._samples_from_blocks(mean[..., -q:], bl, br)

    def samples_minimal(self, X: Tensor) -> Tensor:
        """Same samples from the q new rows only (the algorithm of the CUDA path)."""
        mean_q, S_qb, S_qq = self.model.posterior_minimal_blocks(X, self.X_baseline)
        return self._samples_from_blocks(mean_q, S_qb, S_qq)

    def samples_uncached(self, X: Tensor) -> Tensor:
        """cache_root=False: joint Cholesky of the full (n_b+q) covariance (logei.py:466-472)."""
        q = X.shape[-2]
        Xb = self.X_baseline.expand(*X.shape[:-2], *self.X_baseline.shape[-2:])
        mean, cov = self.model.posterior(torch.cat([Xb, X], dim=-2))
        root = psd_safe_cholesky(cov)
        Z = torch.cat([self.Z_b, self.Z_q(q)], dim=-1)
        full = torch.einsum("...ij,sj->s...i", root, Z) + mean
        return full[..., -q:], full[..., :-q]

    # -- acquisition value -------------------------------------------------------------------
    def _reduce(self, obj: Tensor, best_f: Tensor) -> Tensor:
        # obj: S x B x (M?) x q ; best_f: S x (M?) -> S x 1 x (M?)   (logei.py:422-446)
        bf = best_f.unsqueeze(1)
        acq = safe_math.qlognei_reduction(obj, bf, self.tau_relu, self.tau_max, fat=self.fat)  # B x (M?)
        if self.is_ensemble:  # transforms.py:289-293
            acq = safe_math.logmeanexp(acq, dim=-1)
        return acq

    def forward(self, X: Tensor, mode: str = "faithful") -> Tensor:
        """X: B x q x d -> B."""
        if X.dim() == 2:
            X = X.unsqueeze(0)
        if mode == "faithful":
            obj = self.samples_faithful(X)
        elif mode == "minimal":
            obj = self.samples_minimal(X)
        elif mode == "uncached":
            obj, base = self.samples_uncached(X)
            return self._reduce_uncached(obj, base)
        elif mode == "fallback":
            # cache_root=True whose cached update raised NanError / NotPSDError (cached_cholesky.py:139-165): standard
            # sampling from the joint posterior, last q rows, while best_f stays the cached `_baseline_best_f`
            obj, _ = self.samples_uncached(X)
        else:
            raise ValueError(mode)
        return self._reduce(self.obj_weight * obj, self.best_f)

    def _reduce_uncached(self, obj: Tensor, base: Tensor) -> Tensor:
        obj, base = self.obj_weight * obj, self.obj_weight * base
        best_f = base.max(dim=-1).values  # S x B x (M?)
        li = safe_math.log_improvement(obj, best_f, tau=self.tau_relu)
        acq = safe_math.logmeanexp(safe_math.fatmax(li, dim=-1, tau=self.tau_max), dim=0)
        if self.is_ensemble:
            acq = safe_math.logmeanexp(acq, dim=-1)
        return acq

    def value_and_grad(self, X: Tensor, mode: str = "faithful"):
        """(acq: B, d acq.sum() / dX: B x q x d) as gen_candidates_scipy obtains them (gen.py:201-211)."""
        Xg = X.detach().clone().requires_grad_(True)
        acq = self.forward(Xg, mode=mode)
        (g,) = torch.autograd.grad(acq.sum(), Xg)
        return acq.detach(), g
