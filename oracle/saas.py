"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the potential energy the reference's NUTS fit works on
(``fit_fully_bayesian_model_nuts`` botorch/fit.py:335-396 over ``SaasPyroModel.sample``
models/fully_bayesian.py:167-246).  Only tests/ and tools/ benchmarks may import this module.

pyro (not installed here; pinned by the reference's requirements as pyro-ppl>=1.8.4) turns the model into
``potential_fn(u) = -( sum_sites log p(x_site) + sum_sites log|d x_site / d u_site| )`` with ``x = biject_to(support)(u)``
(pyro.infer.mcmc.util.initialize_model / _PEMaker): the published construction, restated here with torch.distributions'
own ``biject_to`` so that nothing about the transforms is hand-derived.  Sites (models/fully_bayesian.py):
  outputscale ~ Gamma(2, 0.15)                                  :192-203
  mean        ~ Normal(0, 1)                                    :205-214
  noise       = MIN_INFERRED_NOISE_LEVEL + Gamma(0.9, 10)       :216-227   (MIN_INFERRED_NOISE_LEVEL = 1e-4, gpytorch_modules.py:29)
  kernel_tausq ~ HalfCauchy(0.1), _kernel_inv_length_sq ~ HalfCauchy(1)^d, lengthscale = (tausq * inv_length_sq)^-1/2   :229-246
  Y ~ MVN(mean 1, outputscale * matern52(X / lengthscale) + noise I)       :178-190, kernel :81-91
PARITY UNPINNED by reference fixtures (pyro and gpytorch are absent): pinned through torch.distributions only.

Flat layout of the unconstrained vector u (length d + 4), shared with botorch_b200/fully_bayesian.py:
  [ kernel_tausq, _kernel_inv_length_sq (d), mean, noise, outputscale ]
"""
import math
from typing import Dict, Tuple

import torch
from torch import Tensor
from torch.distributions import Gamma, HalfCauchy, MultivariateNormal, Normal, biject_to

MIN_INFERRED_NOISE_LEVEL = 1e-4


def matern52_kernel(X: Tensor, lengthscale: Tensor) -> Tensor:
    """models/fully_bayesian.py:81-91 (``dist`` = gpytorch's euclidean distance: sqrt(clamp_min(sq_dist, 1e-30)))."""
    Z = X / lengthscale
    diff = Z.unsqueeze(-2) - Z.unsqueeze(-3)
    r = diff.pow(2).sum(-1).clamp_min(1e-30).sqrt()
    s5 = math.sqrt(5.0) * r
    return (s5 + (1.0 + 5.0 / 3.0 * r * r)) * torch.exp(-s5)


def sites(d: int, dtype=torch.float64) -> Dict[str, torch.distributions.Distribution]:
    t = lambda v: torch.tensor(v, dtype=dtype)
    return {
        "kernel_tausq": HalfCauchy(t(0.1)),
        "_kernel_inv_length_sq": HalfCauchy(torch.ones(d, dtype=dtype)),
        "mean": Normal(t(0.0), t(1.0)),
        "noise": Gamma(t(0.9), t(10.0)),
        "outputscale": Gamma(t(2.0), t(0.15)),
    }


def split(u: Tensor, d: int) -> Dict[str, Tensor]:
    return {"kernel_tausq": u[0], "_kernel_inv_length_sq": u[1 : 1 + d], "mean": u[1 + d], "noise": u[2 + d],
            "outputscale": u[3 + d]}  # fmt: skip


def potential(u: Tensor, X: Tensor, y: Tensor) -> Tensor:
    """Potential energy of ONE unconstrained vector u [d+4] (differentiable)."""
    d = X.shape[-1]
    parts = split(u, d)
    logp = u.new_zeros(())
    x = {}
    for name, dist in sites(d, u.dtype).items():
        tr = biject_to(dist.support)
        x[name] = tr(parts[name])
        logp = logp + dist.log_prob(x[name]).sum() + tr.log_abs_det_jacobian(parts[name], x[name]).sum()
    lengthscale = (x["kernel_tausq"] * x["_kernel_inv_length_sq"]).rsqrt()
    noise = MIN_INFERRED_NOISE_LEVEL + x["noise"]
    n = X.shape[0]
    K = x["outputscale"] * matern52_kernel(X, lengthscale) + noise * torch.eye(n, dtype=u.dtype)
    logp = logp + MultivariateNormal(loc=x["mean"].view(-1).expand(n), covariance_matrix=K).log_prob(y.reshape(-1))
    return -logp


def potential_and_grad(U: Tensor, X: Tensor, y: Tensor, divergent_as_inf: bool = False) -> Tuple[Tensor, Tensor]:
    """U: [M x (d+4)] -> (energy [M], gradient [M x (d+4)]) by autograd, one chain at a time.  With ``divergent_as_inf`` a
    chain whose evaluation raises (LinAlgError / "PositiveDefinite" ValueError: the exceptions the reference registers with
    pyro's integrator, fully_bayesian.py:67-78) or is not finite gets energy +inf and a zero gradient."""
    vals, grads = [], []
    for u in U:
        ug = u.detach().clone().requires_grad_(True)
        try:
            e = potential(ug, X, y)
            (g,) = torch.autograd.grad(e, ug)
            if divergent_as_inf and not (torch.isfinite(e) and torch.isfinite(g).all()):
                raise ValueError("not finite")
        except (ValueError, torch.linalg.LinAlgError):
            if not divergent_as_inf:
                raise
            e, g = torch.tensor(float("inf"), dtype=u.dtype), torch.zeros_like(u)
        vals.append(e.detach())
        grads.append(g)
    return torch.stack(vals), torch.stack(grads)


def constrained(U: Tensor, d: int) -> Dict[str, Tensor]:
    """What ``postprocess_mcmc_samples`` hands to ``load_mcmc_samples`` (:248-265): mean, noise, outputscale, lengthscale."""
    tausq, ils = U[:, 0].exp(), U[:, 1 : 1 + d].exp()
    # "noise" is the raw pyro site (the Gamma draw): the likelihood adds MIN_INFERRED_NOISE_LEVEL inside `sample_noise`
    # (:215-221) but MCMC records the site value, and `load_mcmc_samples` clamps it from below (:298)
    return {"mean": U[:, 1 + d], "noise": U[:, 2 + d].exp(), "outputscale": U[:, 3 + d].exp(),
            "lengthscale": (tausq.unsqueeze(-1) * ils).rsqrt()}  # fmt: skip
