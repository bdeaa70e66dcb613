"""Oracle restatement of the other improvement-based MC acquisition functions (TEST INFRASTRUCTURE, float64 torch CPU).

Follows
  * /root/reference/botorch/acquisition/logei.py:138-234        qLogExpectedImprovement
  * /root/reference/botorch/acquisition/monte_carlo.py:333-414  qExpectedImprovement
  * /root/reference/botorch/acquisition/monte_carlo.py:417-620  qNoisyExpectedImprovement
  * /root/reference/botorch/acquisition/monte_carlo.py:253-289  forward: sample_reduction(q_reduction(_sample_forward(obj)))
  * /root/reference/botorch/posteriors/gpytorch.py:121-165      rsample_from_base_samples = mean + chol(Sigma) z
  * /root/reference/botorch/utils/transforms.py:289-293         ensemble reduction: mean (plain) / logmeanexp (log)
"""
from __future__ import annotations

import torch
from torch import Tensor

from . import qmc, safe_math
from .gp import OracleGP, psd_safe_cholesky
from .qlognei import TAU_MAX, TAU_RELU, OracleQLogNEI


def _plain_reduction(obj: Tensor, best_f: Tensor) -> Tensor:
    """(obj - best_f).clamp_min(0) -> amax over q -> mean over samples (monte_carlo.py:410, 180-252)."""
    return (obj - best_f.unsqueeze(-1)).clamp_min(0).amax(dim=-1).mean(dim=0)


class OracleQNEI(OracleQLogNEI):
    """qNoisyExpectedImprovement: the qLogNEI sample path with the plain reductions."""

    def _reduce(self, obj: Tensor, best_f: Tensor) -> Tensor:
        acq = _plain_reduction(obj, best_f.unsqueeze(1))
        return acq.mean(dim=-1) if self.is_ensemble else acq


class _OracleEIBase:
    """Posterior over the q candidate points only, sampled through its own q x q Cholesky."""

    log = True

    def __init__(self, model: OracleGP, best_f, num_samples: int = 512, seed: int = 1234, tau_relu: float = TAU_RELU,
                 tau_max: float = TAU_MAX, obj_weight: float = 1.0):  # fmt: skip
        self.model = model
        self.obj_weight = float(obj_weight)
        self.best_f = torch.as_tensor(best_f, dtype=torch.float64)
        self.S, self.seed = num_samples, seed
        self.tau_relu, self.tau_max = tau_relu, tau_max
        self.is_ensemble = len(model.hyper.batch_shape) > 0

    def Z_q(self, q: int) -> Tensor:
        # base samples S x 1 x q, collapsed over the t-batch and the model batch (sampling/base.py:78-96)
        return qmc.sobol_normal(d=q, n=self.S, seed=self.seed)

    def samples(self, X: Tensor) -> Tensor:
        mean, cov = self.model.posterior(X)  # B x (M?) x q, B x (M?) x q x q
        root = psd_safe_cholesky(cov)
        return torch.einsum("...ij,sj->s...i", root, self.Z_q(X.shape[-2])) + mean

    def forward(self, X: Tensor) -> Tensor:
        if X.dim() == 2:
            X = X.unsqueeze(0)
        obj = self.obj_weight * self.samples(X)  # S x B x (M?) x q
        bf = self.best_f  # scalar or (M)
        if self.log:
            li = safe_math.log_improvement(obj, bf, tau=self.tau_relu)
            acq = safe_math.logmeanexp(safe_math.fatmax(li, dim=-1, tau=self.tau_max), dim=0)
            return safe_math.logmeanexp(acq, dim=-1) if self.is_ensemble else acq
        acq = _plain_reduction(obj, bf)
        return acq.mean(dim=-1) if self.is_ensemble else acq

    def value_and_grad(self, X: Tensor):
        Xg = X.detach().clone().requires_grad_(True)
        acq = self.forward(Xg)
        (g,) = torch.autograd.grad(acq.sum(), Xg)
        return acq.detach(), g


class OracleQLogEI(_OracleEIBase):
    log = True


class OracleQEI(_OracleEIBase):
    log = False
