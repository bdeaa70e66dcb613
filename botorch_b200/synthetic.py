"""Synthetic GP problems of the BASELINE.json configs (host-side, float64).

Data generators only -- no model arithmetic.  The three test functions restate
/root/reference/botorch/test_functions/synthetic.py (Ackley :121-167, Branin :181-206,
Hartmann :359-454) and are pinned to it by tests/golden/synthetic_functions.pt.
Input/hyperparameter recipe: SURVEY.md section 8(d) "Synthetic inputs".
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, Optional

import torch
from torch import Tensor
from torch.quasirandom import SobolEngine

_H6_A = [
    [10, 3, 17, 3.5, 1.7, 8],
    [0.05, 10, 17, 0.1, 8, 14],
    [3, 3.5, 1.7, 10, 17, 8],
    [17, 8, 0.05, 10, 0.1, 14],
]
_H6_P = [
    [1312, 1696, 5569, 124, 8283, 5886],
    [2329, 4135, 8307, 3736, 1004, 9991],
    [2348, 1451, 3522, 2883, 3047, 6650],
    [4047, 8828, 8732, 5743, 1091, 381.0],
]
_H_ALPHA = [1.0, 1.2, 3.0, 3.2]


def branin(X: Tensor) -> Tensor:
    """Branin on [-5,10]x[0,15] (synthetic.py:196-205)."""
    x0, x1 = X[..., 0], X[..., 1]
    t1 = x1 - 5.1 / (4 * math.pi**2) * x0.pow(2) + 5 / math.pi * x0 - 6
    t2 = 10 * (1 - 1 / (8 * math.pi)) * torch.cos(x0)
    return t1.pow(2) + t2 + 10


def hartmann6(X: Tensor) -> Tensor:
    """Hartmann-6 on [0,1]^6 (synthetic.py:444-452). float32 constants as in the reference buffers."""
    A = torch.tensor(_H6_A).to(X)  # reference registers float32 buffers then casts (.to(dtype=X.dtype))
    P = torch.tensor(_H6_P).to(X)
    alpha = torch.tensor(_H_ALPHA).to(X)
    inner = torch.sum(A * (X.unsqueeze(-2) - 0.0001 * P).pow(2), dim=-1)
    return -(torch.sum(alpha * torch.exp(-inner), dim=-1))


def ackley(X: Tensor) -> Tensor:
    """Ackley on [-32.768, 32.768]^d (synthetic.py:162-166)."""
    d = X.shape[-1]
    a, b, c = 20, 0.2, 2 * math.pi
    part1 = -a * torch.exp(-b / math.sqrt(d) * torch.linalg.norm(X, dim=-1))
    part2 = -(torch.exp(torch.mean(torch.cos(c * X), dim=-1)))
    return part1 + part2 + a + math.e


_BOUNDS = {
    "branin": [(-5.0, 10.0), (0.0, 15.0)],
    "hartmann6": [(0.0, 1.0)] * 6,
}


def default_lengthscale(d: int) -> float:
    """Prior mode of LogNormal(sqrt2 + log(d)/2, sqrt3) (gpytorch_modules.py:118), floored at 0.025."""
    return max(math.exp(math.sqrt(2) + 0.5 * math.log(d) - 3.0), 0.025)


@dataclass
class Problem:
    """Everything a config needs: training data on the unit cube + fixed hyperparameters."""

    name: str
    train_X: Tensor  # n x d in [0,1]^d
    train_Y: Tensor  # n x 1 raw outcomes (Standardize is applied by the model)
    kernel: str  # "rbf" | "matern52"
    hypers: Dict[str, Tensor]  # lengthscale (d | M x d), noise, mean, [outputscale]
    q: int
    mc_samples: int
    num_restarts: int
    raw_samples: int
    standardize: bool = True
    extra: Dict[str, object] = field(default_factory=dict)

    @property
    def n(self) -> int:
        return self.train_X.shape[0]

    @property
    def d(self) -> int:
        return self.train_X.shape[1]


def _sobol_unit(n: int, d: int, seed: int) -> Tensor:
    return SobolEngine(d, scramble=True, seed=seed).draw(n, dtype=torch.float64)


def _single_task_hypers(d: int, noise: float) -> Dict[str, Tensor]:
    return {
        "lengthscale": torch.full((d,), default_lengthscale(d), dtype=torch.float64),
        "noise": torch.tensor(noise, dtype=torch.float64),
        "mean": torch.tensor(0.0, dtype=torch.float64),
    }


def make_problem(config: str, noise: float = 1e-2, n: Optional[int] = None, M: Optional[int] = None) -> Problem:
    """config in {"C1","C2","C3","C4","C5"}. ``n``/``M`` shrink a config for tests."""
    config = config.upper()
    if config == "C1":
        n = n or 20
        X = _sobol_unit(n, 2, 0)
        lo = torch.tensor([b[0] for b in _BOUNDS["branin"]], dtype=torch.float64)
        hi = torch.tensor([b[1] for b in _BOUNDS["branin"]], dtype=torch.float64)
        Y = -branin(lo + (hi - lo) * X)
        return Problem("C1-branin", X, Y.unsqueeze(-1), "rbf", _single_task_hypers(2, noise), 1, 512, 10, 512)
    if config == "C2":
        n = n or 200
        X = _sobol_unit(n, 6, 0)
        Y = -hartmann6(X)
        return Problem("C2-hartmann6", X, Y.unsqueeze(-1), "rbf", _single_task_hypers(6, noise), 4, 256, 64, 4096)
    if config == "C3":
        n = n or 2000
        X = _sobol_unit(n, 20, 0)
        Y = -ackley(-32.768 + 65.536 * X)
        return Problem("C3-ackley20", X, Y.unsqueeze(-1), "rbf", _single_task_hypers(20, noise), 8, 256, 512, 32768)
    if config == "C4":
        n, M, d = n or 100, M or 256, 30
        g = torch.Generator().manual_seed(0)
        X = _sobol_unit(n, d, 0)
        Y = torch.sin(2 * math.pi * X[:, 0]) + 0.1 * torch.randn(n, generator=g, dtype=torch.float64)
        # SAAS-like draws: most dimensions have long lengthscales, a few are short
        ls = 0.1 + 1.9 * torch.rand(M, d, generator=g, dtype=torch.float64)
        ls[:, 3:] = ls[:, 3:] * 5.0
        hypers = {
            "lengthscale": ls,
            "outputscale": 0.5 + 1.5 * torch.rand(M, generator=g, dtype=torch.float64),
            "noise": 1e-3 + (1e-1 - 1e-3) * torch.rand(M, generator=g, dtype=torch.float64),
            "mean": 0.1 * torch.randn(M, generator=g, dtype=torch.float64),
        }
        return Problem("C4-saas", X, Y.unsqueeze(-1), "matern52", hypers, 4, 256, 64, 4096, standardize=False)
    if config == "C5":
        n, d = n or 20000, 12
        g = torch.Generator().manual_seed(0)
        X = _sobol_unit(n, d, 0)
        Y = torch.sin(2 * math.pi * X[:, 0]) + 0.1 * torch.randn(n, generator=g, dtype=torch.float64)
        p = Problem("C5-thompson", X, Y.unsqueeze(-1), "rbf", _single_task_hypers(d, noise), 1, 1, 0, 0)
        p.extra["num_candidates"] = 1_000_000
        return p
    raise ValueError(f"unknown config {config}")
