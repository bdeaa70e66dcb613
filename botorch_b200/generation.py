"""Thompson sampling over a discrete candidate set (config C5).

Contract mirrored from ``MaxPosteriorSampling`` (botorch/generation/sampling.py:48-155): ``forward(X, num_samples)``
returns the ``num_samples`` candidates maximising posterior draws.

The reference draws from the JOINT posterior over all N candidates (an N x N Cholesky; the tutorial caps N at
10 000).  Up to ``joint_max`` candidates this class does the same on the GPU: GEMM-shaped joint posterior
(``bo_posterior_joint_large``), blocked Cholesky, one triangular GEMM for the draws, then the reference's
``maximize_samples`` index logic.  At the scale BASELINE.json names (N = 1e6, n = 20 000) only the marginal form
exists: each candidate is its own t-batch (``X.unsqueeze(-2)``, q = 1), i.e. mean_i + sqrt(var_i) z_i -- the
cross-covariance GEMM + triangular-solve path of SURVEY.md row T1; there ``replacement=False`` removes already picked
candidates draw by draw.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
from torch import Tensor
from torch.nn import Module

from . import _lib
from .exceptions import UnsupportedError
from .models import _FusedExactGP


def _flip_sub_unique(x: Tensor, k: int) -> Tensor:
    """First k unique elements of a 1-d tensor, traversed from the back (generation/utils.py:51-82)."""
    n = len(x)
    seen, keep = set(), []
    for j, xi in enumerate(x.flip(0).tolist()):
        if xi not in seen:
            seen.add(xi)
            keep.append(n - 1 - j)
        if len(seen) >= k:
            break
    return x[torch.tensor(keep, dtype=torch.long, device=x.device)]


class MaxPosteriorSampling(Module):
    def __init__(self, model, objective=None, posterior_transform=None, replacement: bool = True,
                 chunk_size: int = 16384, joint_max: int = 16384) -> None:  # fmt: skip
        super().__init__()
        if objective is not None or posterior_transform is not None:
            raise UnsupportedError("objective / posterior_transform are outside the fused path")
        if not isinstance(model, _FusedExactGP):
            raise UnsupportedError("MaxPosteriorSampling on the fused path needs a botorch_b200 exact GP model")
        self.model = model
        self.replacement = replacement
        self.chunk_size = int(chunk_size)
        self.joint_max = int(joint_max)  # largest N sampled jointly (N x N covariance: 2 GiB at 16384)

    def marginal_posterior(self, X: Tensor):
        """mean, var [M x N] of N candidates, streamed through the GEMM pipeline in chunks."""
        st = self.model.gp_state
        N = X.shape[0]
        mean = torch.empty(st.M, N, dtype=torch.float64, device=st.device)
        var = torch.empty(st.M, N, dtype=torch.float64, device=st.device)
        work = st.posterior_workspace(min(self.chunk_size, N))
        for i in range(0, N, self.chunk_size):
            m, v = st.posterior_mean_var(X[i : i + self.chunk_size], work=work)
            mean[:, i : i + self.chunk_size], var[:, i : i + self.chunk_size] = m, v
        return mean, var

    def maximize_samples(self, X: Tensor, samples: Tensor, num_samples: int = 1) -> Tensor:
        """sampling.py:112-155 for one candidate set: samples [num_samples x (M) x N x 1] -> picked rows of X."""
        obj = samples.squeeze(-1)  # IdentityMCObjective
        if self.replacement:
            idcs = torch.argmax(obj, dim=-1)
        else:
            _, idcs_full = torch.topk(obj, num_samples, dim=-1)
            ridx, cindx = torch.tril_indices(num_samples, num_samples)
            sub_idcs = idcs_full[ridx, ..., cindx]
            if sub_idcs.ndim == 1:
                idcs = _flip_sub_unique(sub_idcs, num_samples)
            elif sub_idcs.ndim == 2:
                idcs = torch.stack([_flip_sub_unique(sub_idcs[:, i], num_samples) for i in range(sub_idcs.size(-1))], dim=-1)
            else:
                raise NotImplementedError(
                    "MaxPosteriorSampling without replacement for more than a single batch dimension is not yet implemented."
                )
        if idcs.ndim > 1:
            idcs = idcs.permute(*range(1, idcs.ndim), 0)
        idcs = idcs.unsqueeze(-1).expand(*idcs.shape, X.size(-1))
        Xe = X.expand(*obj.shape[1:], X.size(-1))
        return torch.gather(Xe, -2, idcs)

    @torch.no_grad()
    def forward(self, X: Tensor, num_samples: int = 1, observation_noise: bool = False,
                generator: Optional[torch.Generator] = None, joint: Optional[bool] = None) -> Tensor:  # fmt: skip
        """X: N x d -> num_samples x d maximisers of posterior draws: joint draws over all N candidates (the
        reference's semantics) when ``N <= joint_max`` (override with ``joint``), independent marginal draws beyond."""
        if X.dim() != 2:
            raise UnsupportedError("the fused Thompson sampler takes one N x d candidate set")
        st = self.model.gp_state
        if X.device != st.device:
            raise _lib.BoError(f"X is on {X.device}; the model lives on {st.device}. There is no CPU path.")
        if joint is None:
            joint = X.shape[0] <= self.joint_max
        if joint:
            posterior = self.model.posterior(X, observation_noise=observation_noise)
            samples = posterior.rsample(sample_shape=torch.Size([num_samples]))
            return self.maximize_samples(X, samples, num_samples)
        Xd = X.to(torch.float64).contiguous()
        mean, var = self.marginal_posterior(Xd)
        if observation_noise:
            var = var + (st.y_std**2) * st.noise.unsqueeze(-1)
        if st.M > 1:  # mixture over hyperparameter samples: each draw uses one component, like sampling the mixture
            comp = torch.randint(st.M, (num_samples,), device=st.device, generator=generator)
        else:
            comp = torch.zeros(num_samples, dtype=torch.long, device=st.device)
        N = Xd.shape[0]
        lib = _lib.lib()
        picked = []
        val = torch.empty(1, dtype=torch.float64, device=st.device)
        idx = torch.empty(1, dtype=torch.int64, device=st.device)
        mean_s = mean.clone() if not self.replacement else mean
        for s in range(num_samples):
            z = torch.randn(N, dtype=torch.float64, device=st.device, generator=generator)
            m = int(comp[s]) if st.M > 1 else 0
            _lib.check(lib.bo_ts_argmax(st.ctx, 1, N, mean_s[m].data_ptr(), var[m].data_ptr(), z.data_ptr(),
                                        val.data_ptr(), idx.data_ptr(), _lib.stream_ptr(st.device)))  # fmt: skip
            picked.append(idx.clone())
            if not self.replacement:
                mean_s[:, idx] = -float("inf")
        return X[torch.cat(picked)]
