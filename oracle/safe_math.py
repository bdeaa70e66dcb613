"""Oracle restatement of the qLogNEI reductions (TEST INFRASTRUCTURE).

Follows /root/reference/botorch/utils/safe_math.py:
  * ``log_fatplus`` / ``fatplus`` / ``cauchy``      -> safe_math.py:293-320, 449-451
  * ``fatmax`` / ``_pareto`` / ``_inf_max_helper``   -> safe_math.py:323-352, 454-478, 149-187
  * ``logsumexp`` / ``logmeanexp``                   -> safe_math.py:124-146, 209-223
and ``_log_improvement`` of /root/reference/botorch/acquisition/logei.py:509-534.

All functions are torch float64 and differentiable, so ``torch.autograd`` on the
oracle gives the gradient the reference's autograd would give (the reference has
no hand-written backward on this path).  Pinned against the verbatim reference
module by tests/test_oracle_golden.py.
"""
from __future__ import annotations

import math
from typing import Callable, Tuple, Union

import torch
from torch import Tensor
from torch.nn.functional import softplus

Dim = Union[int, Tuple[int, ...]]

FATPLUS_CAUCHY_WEIGHT = 0.1  # safe_math.py:316 (alpha of `_fatplus`)


def lorentzian(x: Tensor) -> Tensor:
    """Un-normalised Cauchy density 1/(1+x^2)  (safe_math.py:449-451)."""
    return 1 / (1 + x.square())


def fatplus(x: Tensor, tau: float) -> Tensor:
    """tau*(softplus(x/tau) + 0.1*cauchy(x/tau))  (safe_math.py:302-320).

    ``softplus`` is torch's (beta=1, threshold=20): identity above 20.
    """
    u = x / tau
    return tau * (softplus(u) + FATPLUS_CAUCHY_WEIGHT * lorentzian(u))


def log_fatplus(x: Tensor, tau: float) -> Tensor:
    """log(fatplus(x, tau))  (safe_math.py:293-299)."""
    return fatplus(x, tau=tau).log()


def log_improvement(Y: Tensor, best_f: Tensor, tau: float) -> Tensor:
    """log_fatplus(Y - best_f[..., None])  (logei.py:509-534, fat=True)."""
    return log_fatplus(Y - best_f.unsqueeze(-1).to(Y), tau=tau)


def pareto(x: Tensor, alpha: float) -> Tensor:
    """Rational fat-tailed decay (safe_math.py:454-478).

    With a = alpha/2: (b0/(b0 + b1*x + x^2))**a, b1 = 2a, b0 = a*b1.
    For the default alpha=2: 2/(2 + 2x + x^2).
    """
    a = alpha / 2
    b1 = 2 * a
    b0 = a * b1
    return (b0 / (b0 + b1 * x + x.square())).pow(a)


def _any_over(x: Tensor, dim: Dim) -> Tensor:
    if isinstance(dim, tuple):
        for d in dim:
            x = x.any(dim=d, keepdim=True)
        return x
    return x.any(dim, keepdim=True)


def inf_aware_max(reduce: Callable[..., Tensor], x: Tensor, dim: Dim, keepdim: bool) -> Tensor:
    """Smooth-max skeleton with explicit handling of infinities (safe_math.py:149-187).

    * slices whose maximum is infinite return the sum of their infinite maxima
      (so +inf stays +inf, an all -inf slice stays -inf, and the gradient is 1
      on those entries);
    * otherwise ``M + reduce(x - M)`` with the finite maximum ``M`` detached
      only through the masked fills exactly as the reference does it.
    """
    M = x.amax(dim=dim, keepdim=True)
    at_inf_max = torch.logical_and(*torch.broadcast_tensors(M.isinf(), x == M))
    slice_has_inf_max = _any_over(at_inf_max, dim)
    only_inf = x.masked_fill(~at_inf_max, 0.0)
    M_fin = M.masked_fill(M.isinf(), 0.0)
    shifted = x.masked_fill(slice_has_inf_max, 0.0) - M_fin
    out = torch.where(
        slice_has_inf_max,
        only_inf.sum(dim=dim, keepdim=True),
        M_fin + reduce(shifted, dim=dim, keepdim=True),
    )
    return out if keepdim else out.sum(dim=dim)


def logsumexp(x: Tensor, dim: Dim, keepdim: bool = False) -> Tensor:
    """safe_math.py:124-146."""
    return inf_aware_max(torch.logsumexp, x=x, dim=dim, keepdim=keepdim)


def logmeanexp(X: Tensor, dim: Dim, keepdim: bool = False) -> Tensor:
    """safe_math.py:209-223."""
    n = X.shape[dim] if isinstance(dim, int) else math.prod(X.shape[i] for i in dim)
    return logsumexp(X, dim=dim, keepdim=keepdim) - math.log(n)


def fatmax(x: Tensor, dim: Dim, keepdim: bool = False, tau: float = 1.0, alpha: float = 2.0) -> Tensor:
    """Fat-tailed smooth maximum (safe_math.py:323-352)."""

    def reduce(y: Tensor, dim: Dim, keepdim: bool = False) -> Tensor:
        return tau * pareto(-y / tau, alpha=alpha).sum(dim=dim, keepdim=keepdim).log()

    return inf_aware_max(reduce, x=x, dim=dim, keepdim=keepdim)


def log_softplus(x: Tensor, tau: float) -> Tensor:
    """log(softplus(x, beta=1/tau)) with the reference's cut-offs (safe_math.py:226-247; float64: upper 32, lower -35)."""
    tau_t = torch.as_tensor(tau, dtype=x.dtype)
    upper, lower = (16, -15) if x.dtype == torch.float32 else (32, -35)
    mask = x / tau_t > lower
    return torch.where(
        mask,
        softplus(x.masked_fill(~mask, lower), beta=(1 / tau_t), threshold=upper).log(),
        x / tau_t + tau_t.log(),
    )


def smooth_amax(x: Tensor, dim: Dim = -1, keepdim: bool = False, tau: float = 1.0) -> Tensor:
    """tau * logsumexp(x / tau)  (safe_math.py:250-273)."""
    return logsumexp(x / tau, dim=dim, keepdim=keepdim) * tau


def qlognei_reduction(obj: Tensor, best_f: Tensor, tau_relu: float, tau_max: float, fat: bool = True) -> Tensor:
    """samples -> acquisition value (monte_carlo.py:253-274 with logei.py:122,129,347-362).

    obj: S x batch x q objective samples, best_f: S (x model batch) best baseline
    objective per MC sample.  Returns ``batch``.
    """
    if fat:
        li = log_improvement(obj, best_f, tau=tau_relu)
        per_sample = fatmax(li, dim=-1, tau=tau_max)
    else:  # logei.py:122 (smooth_amax), :532 (log_softplus)
        li = log_softplus(obj - best_f.unsqueeze(-1).to(obj), tau=tau_relu)
        per_sample = smooth_amax(li, dim=-1, tau=tau_max)
    return logmeanexp(per_sample, dim=0)
