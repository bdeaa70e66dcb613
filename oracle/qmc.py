"""Oracle restatement of the base-sample / raw-candidate draws (TEST INFRASTRUCTURE).

Follows
  * /root/reference/botorch/sampling/qmc.py:59-101   NormalQMCEngine.draw (inv_transform=True)
  * /root/reference/botorch/utils/sampling.py:64-136 draw_sobol_samples / draw_sobol_normal_samples
  * /root/reference/botorch/sampling/normal.py:178-209 SobolQMCNormalSampler._construct_base_samples
  * /root/reference/botorch/sampling/normal.py:68-131  _update_base_samples (baseline columns re-used)
The bit generator is torch's own ``SobolEngine`` (host, as in the reference), so
the draws are bit-identical to the reference by construction; the golden test
checks that against fixtures produced by the verbatim reference module.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import Tensor
from torch.quasirandom import SobolEngine


def sobol_normal(d: int, n: int, seed: Optional[int], dtype=torch.float64) -> Tensor:
    """n x d scrambled-Sobol N(0,1) draws through the inverse CDF (qmc.py:79-84)."""
    u = SobolEngine(dimension=d, scramble=True, seed=seed).draw(n, dtype=dtype)
    v = 0.5 + (1 - torch.finfo(u.dtype).eps) * (u - 0.5)
    return torch.erfinv(2 * v - 1) * math.sqrt(2)


def sobol_box(bounds: Tensor, n: int, q: int, seed: Optional[int]) -> Tensor:
    """n x q x d scrambled-Sobol q-batches in the box ``bounds`` (sampling.py:64-104)."""
    d = bounds.shape[-1]
    raw = SobolEngine(q * d, scramble=True, seed=seed).draw(n, dtype=bounds.dtype)
    return bounds[0] + (bounds[1] - bounds[0]) * raw.view(n, q, d)


def qlognei_base_samples(S: int, n_b: int, q: int, seed: int, dtype=torch.float64):
    """Base samples exactly as qLogNEI with cache_root=True ends up holding them.

    ``_init_baseline`` draws S x n_b samples for the baseline posterior
    (logei.py:395-413; sampler seed ``seed``); on the first forward with q new
    points ``_set_sampler`` -> ``_update_base_samples`` (cached_cholesky.py:167-186,
    normal.py:68-131) draws a fresh S x (n_b+q) set with the SAME seed and then
    overwrites its first n_b columns with the baseline draw.

    Returns (Z_b: S x n_b, Z_q: S x q).
    """
    z_base = sobol_normal(d=n_b, n=S, seed=seed, dtype=dtype)
    z_full = sobol_normal(d=n_b + q, n=S, seed=seed, dtype=dtype)
    return z_base, z_full[:, n_b:].clone()
