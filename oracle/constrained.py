"""Oracle restatement of constrained qLogNEI over independent single-output models (TEST INFRASTRUCTURE, float64 CPU).

  reference                                                        restated here
  ---------------------------------------------------------------  ----------------------------------------------
  acquisition/monte_carlo.py:253-330 (forward, _apply_constraints)   ``OracleConstrainedQLogNEI.forward``
  utils/objective.py:101-180                                         ``feasibility`` / ``smoothed_log_feasibility``
  utils/safe_math.py:97-99, 422-451                                  ``logexpit``, ``fatmoid``, ``log_fatmoid``
  acquisition/utils.py:90-163                                        ``best_feasible_objective`` (no-fallback branch: the
                                                                      tests construct baselines with a feasible point per sample)
  sampling/normal.py:88-131, 178-209; utils/low_rank.py:43-82        base samples S x 1 x (n_b+q) x m, outcome k reads [..., k]
  acquisition/logei.py:364-481                                       per-outcome cached-Cholesky samples (oracle/qlognei.py)

``utils/objective.py`` and ``utils/safe_math.py`` import verbatim from /root/reference: tests/test_oracle_golden.py pins the
restatements against them (and against literals minted from them, for machines without the reference tree).
"""
from __future__ import annotations

import math
from typing import Callable, List, Optional, Sequence

import torch
from torch import Tensor

from . import qmc, safe_math
from .gp import OracleGP, psd_safe_cholesky
from .qlognei import TAU_MAX, TAU_RELU, OracleQLogNEI

_INV_SQRT_3 = math.sqrt(1 / 3)


def logexpit(X: Tensor) -> Tensor:
    return -torch.nn.functional.softplus(-X, threshold=math.log(torch.finfo(X.dtype).max) / 2)


def fatmoid(X: Tensor, tau: float = 1.0) -> Tensor:
    X = X / tau
    cauchy = lambda x: 1 / (1 + x.square())
    return torch.where(X < 0, 2 / 3 * cauchy(X - _INV_SQRT_3), 1 - 2 / 3 * cauchy(X + _INV_SQRT_3))


def log_fatmoid(X: Tensor, tau: float = 1.0) -> Tensor:
    return fatmoid(X, tau=tau).log()


def feasibility(constraints, samples: Tensor) -> Tensor:
    ind = torch.ones(samples.shape[:-1], dtype=torch.bool)
    for c in constraints:
        ind = ind.logical_and(c(samples) <= 0)
    return ind


def smoothed_log_feasibility(constraints, samples: Tensor, eta: Sequence[float], fat: bool = True) -> Tensor:
    out = torch.zeros_like(samples[..., 0])
    f = log_fatmoid if fat else logexpit
    for c, e in zip(constraints, eta):
        out = out + f(-c(samples) / e)
    return out


def prune_inferior_points(models, X: Tensor, constraints, obj_idx: int, obj_w: float, num_samples: int, seed: int,
                          max_frac: float = 1.0) -> Tensor:
    """acquisition/utils.py:245-349 for m independent outcomes: indices (into X) of the points that are the best feasible
    one in at least one joint posterior sample; Sobol base samples over the (points x outcomes) block."""
    n, m = X.shape[0], len(models)
    z = qmc.sobol_normal(d=n * m, n=num_samples, seed=seed).view(num_samples, n, m)
    cols = []
    for k, gp in enumerate(models):
        mean, cov = gp.posterior(X)
        root = psd_safe_cholesky(cov)
        cols.append(mean + torch.einsum("ij,sj->si", root, z[..., k]))
    samples = torch.stack(cols, dim=-1)
    obj = obj_w * samples[..., obj_idx]
    infeas = ~feasibility(constraints, samples)
    if bool(infeas.any()):
        obj[infeas] = obj.min() - 1
    idcs, counts = torch.unique(torch.argmax(obj, dim=-1), return_counts=True)
    if len(idcs) > math.ceil(max_frac * n):
        counts, order = torch.sort(counts, descending=True)
        idcs = order[: math.ceil(max_frac * n)]
    return idcs


def objective_lower_bound(models, X: Tensor, convex_weights: Tensor, obj_idx: int, obj_w: float) -> float:
    """``_estimate_objective_lower_bound`` + ``get_infeasible_cost`` (acquisition/utils.py:166-242): the objective of the
    6-sigma lower posterior bound at convex combinations of the baseline points (``convex_weights``: the 32 x n uniform
    draws, normalised over dim 0 as the reference does), minimum over points, clamped at 0.  Returns -M."""
    w = convex_weights / convex_weights.sum(dim=0, keepdim=True)
    Xc = w @ X
    mean, cov = models[obj_idx].posterior(Xc.unsqueeze(0))  # 1 x 32 joint block; only its diagonal is used
    var = torch.diagonal(cov, dim1=-2, dim2=-1).clamp_min(0)
    lb = obj_w * (mean - 6.0 * var.sqrt())
    return float(lb.min().clamp_max(0.0))


class OracleConstrainedQLogNEI:
    """m independent OracleGPs; objective = obj_weight * outcome obj_idx; constraints: callables on (..., m) samples."""

    def __init__(self, models: List[OracleGP], X_baseline: Tensor, constraints: List[Callable[[Tensor], Tensor]], eta,
                 obj_idx: int = 0, obj_weight: float = 1.0, num_samples: int = 64, seed: int = 1234, fat: bool = True,
                 tau_relu: float = TAU_RELU, tau_max: float = TAU_MAX, convex_weights: Optional[Tensor] = None, plain: bool = False,
                 objective: Optional[Callable[[Tensor, Tensor], Tensor]] = None):  # fmt: skip
        # objective: a GenericMCObjective callable f(samples, X) (acquisition/objective.py:344-384); None = obj_weight * outcome obj_idx
        self.plain, self.objective = plain, objective
        self.m, self.S, self.seed = len(models), num_samples, seed
        self.constraints = constraints
        # utils/objective.py:167-168: a float eta becomes torch.full((n,), eta) in the DEFAULT dtype (float32 unless the caller
        # changed it), so 0.05 enters the division as float32(0.05); a tensor eta is used as given
        self.eta = list(torch.full((len(constraints),), eta)) if not isinstance(eta, Tensor) else list(eta)
        self.fat, self.obj_idx, self.obj_w = fat, obj_idx, float(obj_weight)
        self.tau_relu, self.tau_max = tau_relu, tau_max
        self.X_baseline = X_baseline
        self.n_b = n_b = X_baseline.shape[0]
        Zb = qmc.sobol_normal(d=n_b * self.m, n=self.S, seed=seed).view(self.S, n_b, self.m)
        self.parts = []
        for k, gp in enumerate(models):
            o = OracleQLogNEI(gp, X_baseline, num_samples=self.S, seed=seed, tau_relu=tau_relu, tau_max=tau_max)
            o.Z_b = Zb[:, :, k].contiguous()
            with torch.no_grad():
                mean_b, _ = gp.posterior(X_baseline)
                o.baseline_samples = torch.einsum("ij,sj->si", o.baseline_L, o.Z_b) + mean_b
            self.parts.append(o)
        base = torch.stack([o.baseline_samples for o in self.parts], dim=-1)  # S x n_b x m
        obj_b = self._obj(base, X_baseline)
        feas = feasibility(constraints, base)
        if bool(feas.any(dim=-1).all()):
            infeasible_value = -float("inf")
        else:  # acquisition/utils.py:137-152: no feasible baseline point in some sample -> 6-sigma lower bound
            assert convex_weights is not None, "pass the 32 x n_b uniform draws of _estimate_objective_lower_bound"
            infeasible_value = objective_lower_bound(models, X_baseline, convex_weights, obj_idx, self.obj_w)
        self.best_f = torch.where(feas, obj_b, torch.full_like(obj_b, infeasible_value)).amax(dim=-1)  # S

    def _obj(self, samples: Tensor, X: Tensor) -> Tensor:
        if self.objective is not None:
            return self.objective(samples, X)
        return self.obj_w * samples[..., self.obj_idx]

    def _set_q(self, q: int) -> None:
        full = qmc.sobol_normal(d=(self.n_b + q) * self.m, n=self.S, seed=self.seed).view(self.S, self.n_b + q, self.m)
        for k, o in enumerate(self.parts):
            o._Zq_cache[q] = full[:, self.n_b :, k].contiguous()

    def samples(self, X: Tensor) -> Tensor:
        """S x B x q x m."""
        self._set_q(X.shape[-2])
        return torch.stack([o.samples_faithful(X) for o in self.parts], dim=-1)

    def forward(self, X: Tensor) -> Tensor:
        if X.dim() == 2:
            X = X.unsqueeze(0)
        smp = self.samples(X)
        obj = self._obj(smp, X)
        if self.plain:  # qNEI (monte_carlo.py:305-330, 417-620): relu improvement times the feasibility weights, amax, mean
            imp = (obj - self.best_f.view(self.S, 1, 1)).clamp_min(0)
            if self.constraints:
                imp = imp * smoothed_log_feasibility(self.constraints, smp, self.eta, self.fat).exp()
            return imp.amax(dim=-1).mean(dim=0)
        if self.fat:
            li = safe_math.log_improvement(obj, self.best_f.view(self.S, 1), tau=self.tau_relu)
        else:  # the reference's `fat` flag switches the ReLU smoothing, the q-reduction AND the sigmoid (logei.py:122,532)
            li = safe_math.log_softplus(obj - self.best_f.view(self.S, 1, 1), tau=self.tau_relu)
        if self.constraints:
            li = li + smoothed_log_feasibility(self.constraints, smp, self.eta, self.fat)
        qmax = safe_math.fatmax if self.fat else safe_math.smooth_amax
        return safe_math.logmeanexp(qmax(li, dim=-1, tau=self.tau_max), dim=0)

    def value_and_grad(self, X: Tensor):
        Xg = X.detach().clone().requires_grad_(True)
        acq = self.forward(Xg)
        (g,) = torch.autograd.grad(acq.sum(), Xg)
        return acq.detach(), g


class OracleConstrainedQLogEI(OracleConstrainedQLogNEI):
    """acquisition/logei.py:138-234 (qLogExpectedImprovement) / monte_carlo.py:333-414 (qExpectedImprovement, ``plain``) with
    outcome constraints and/or a generic objective: joint posterior over the q candidate points of every outcome
    (posteriors/gpytorch.py:121-165: mean + chol(Sigma) z), base samples S x 1 x q x m from ONE draw over the q * m block
    (sampling/normal.py:34-86), the FIXED incumbent ``best_f``; ``forward`` is the parent's."""

    def __init__(self, models: List[OracleGP], best_f: float, constraints, eta, obj_idx: int = 0, obj_weight: float = 1.0,
                 num_samples: int = 64, seed: int = 1234, fat: bool = True, tau_relu: float = TAU_RELU,
                 tau_max: float = TAU_MAX, plain: bool = False, objective=None):  # fmt: skip
        self.plain, self.objective = plain, objective
        self.models, self.m, self.S, self.seed = models, len(models), num_samples, seed
        self.constraints = constraints or []
        self.eta = list(torch.full((len(self.constraints),), eta)) if not isinstance(eta, Tensor) else list(eta)
        self.fat, self.obj_idx, self.obj_w = fat, obj_idx, float(obj_weight)
        self.tau_relu, self.tau_max = tau_relu, tau_max
        self.best_f = torch.full((self.S,), float(best_f), dtype=torch.float64)

    def samples(self, X: Tensor) -> Tensor:
        q = X.shape[-2]
        Z = qmc.sobol_normal(d=q * self.m, n=self.S, seed=self.seed).view(self.S, q, self.m)
        cols = []
        for k, gp in enumerate(self.models):
            mean, cov = gp.posterior(X)  # B x q, B x q x q
            cols.append(torch.einsum("...ij,sj->s...i", psd_safe_cholesky(cov), Z[:, :, k]) + mean)
        return torch.stack(cols, dim=-1)  # S x B x q x m
