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

``pathwise=True`` (or ``draw_matheron_paths``) gives JOINT draws at any N: Matheron's rule over random-Fourier-feature
prior paths (botorch/sampling/pathwise/: prior_samplers.py:51-91, update_strategies.py:72-103,
posterior_samplers.py:143-199; acquisition/thompson_sampling.py evaluates such a path on a candidate set).  Each draw
is a function, so the argmax over 1e6 candidates is consistent across them, and a chunk costs one covariance GEMM
k(X, x) plus two skinny GEMMs instead of the n^2-per-candidate triangular product of the marginal variance.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor
from torch.nn import Module

from . import _lib
from .exceptions import UnsupportedError
from .models import _FusedExactGP
from .sampling import draw_sobol_normal_samples


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


class MatheronPath:
    """Sample paths f_s(.) of the GP posterior: ``path(X)`` with X ``(...) x d`` returns ``sample_shape x (...)`` values on
    the outcome scale.  Prior part: F random Fourier features; update part: exact (uses the model's Cholesky factor)."""

    def __init__(self, model: _FusedExactGP, sample_shape: torch.Size, num_features: int = 1024,
                 chunk_size: int = 16384) -> None:  # fmt: skip
        from . import engine

        if model._is_fully_bayesian:
            raise NotImplementedError("PathwiseThompsonSampling is not supported for fully Bayesian models")
        if num_features % 2:
            raise UnsupportedError(f"Expected an even number of output features, but received {num_features=}.")
        model.eval()
        st = model.gp_state
        self.sample_shape = torch.Size(sample_shape)
        S = max(int(self.sample_shape.numel()), 1)
        d, n, half = st.d, st.n, num_features // 2
        # the reference's generators, in its order of draws (features/generators.py:114-155, prior_samplers.py:74-80,
        # update_strategies.py:83): Sobol-normal frequencies and weights, iid normal noise; torch's global generator
        W = draw_sobol_normal_samples(d=d, n=half, dtype=torch.float64)
        if st.kernel == "matern52":
            nu = torch.tensor(2.5, dtype=torch.float64)
            W = torch.distributions.Gamma(nu, nu).rsample((half, 1)).rsqrt() * W
        w = draw_sobol_normal_samples(d=num_features, n=S, dtype=torch.float64)
        eps = torch.randn(S, n, dtype=torch.float64)
        self.state = engine.PathState(st, W.to(st.device), w.to(st.device), eps.to(st.device))
        self.chunk_size = int(chunk_size)

    def __call__(self, X: Tensor) -> Tensor:
        from .engine import path_values

        st = self.state.gp
        if X.device != st.device:
            raise _lib.BoError(f"X is on {X.device}; the model lives on {st.device}. There is no CPU path.")
        lead = X.shape[:-1]
        Xf = X.reshape(-1, X.shape[-1])
        chunks = [path_values(self.state, Xf[i : i + self.chunk_size]) for i in range(0, Xf.shape[0], self.chunk_size)]
        out = chunks[0] if len(chunks) == 1 else torch.cat(chunks, dim=-1)
        return out.reshape(*self.sample_shape, *lead)


def draw_matheron_paths(model: _FusedExactGP, sample_shape: torch.Size, num_features: int = 1024) -> MatheronPath:
    """botorch/sampling/pathwise/posterior_samplers.py:143-169 for the fused exact GP."""
    if not isinstance(model, _FusedExactGP):
        raise UnsupportedError("draw_matheron_paths on the fused path needs a botorch_b200 exact GP model")
    return MatheronPath(model, sample_shape, num_features=num_features)


class PathwiseThompsonSampling(Module):
    """Single-outcome Thompson sampling via sample paths (botorch/acquisition/thompson_sampling.py:30-87): the q points of a
    candidate batch are scored by q independent posterior function draws, ``forward(X: b x q x d) -> b`` sums
    ``f_j(x_j)`` over the batch, differentiable w.r.t. X so ``optimize_acqf`` can maximise it.  The paths are drawn on the
    first call (for that q) and kept until ``redraw()``."""

    def __init__(self, model: _FusedExactGP, posterior_transform=None, num_features: int = 1024) -> None:
        super().__init__()
        if posterior_transform is not None:
            raise UnsupportedError("posterior_transform is outside the fused path")
        if model._is_fully_bayesian:
            raise NotImplementedError("PathwiseThompsonSampling is not supported for fully Bayesian models")
        if not isinstance(model, _FusedExactGP):
            raise UnsupportedError("PathwiseThompsonSampling on the fused path needs a botorch_b200 exact GP model")
        self.model = model
        self.num_features = int(num_features)
        self.batch_size: Optional[int] = None
        self.samples: Optional[MatheronPath] = None

    def set_X_pending(self, X_pending=None) -> None:
        if X_pending is not None:
            raise UnsupportedError("X_pending is not used by pathwise Thompson sampling")

    def redraw(self) -> None:
        self.samples = MatheronPath(self.model, torch.Size([self.batch_size]), self.num_features)

    def forward(self, X: Tensor) -> Tensor:
        if X.dim() < 2:
            raise ValueError(f"Expected X to be `(b) x q x d`, got {tuple(X.shape)}")
        X3 = X if X.dim() > 2 else X.unsqueeze(0)
        q = X3.shape[-2]
        if self.batch_size is None:
            self.batch_size = q
            self.redraw()
        elif self.batch_size != q:
            raise ValueError(
                f"The batch size of PathwiseThompsonSampling should not change during a forward pass - was {self.batch_size}, "
                f"now {q}. Please re-initialize the acquisition if you want to change the batch size."
            )
        vals = self.samples(X3)  # [q paths] x batch_shape x [q points]
        own = torch.diagonal(vals.movedim(0, -2), dim1=-2, dim2=-1)  # path j at point j
        return own.sum(dim=-1)


class MaxPosteriorSampling(Module):
    def __init__(self, model, objective=None, posterior_transform=None, replacement: bool = True,
                 chunk_size: int = 16384, joint_max: int = 16384, pathwise: bool = False,
                 num_features: int = 1024) -> None:  # fmt: skip
        super().__init__()
        if objective is not None or posterior_transform is not None:
            raise UnsupportedError("objective / posterior_transform are outside the fused path")
        if not isinstance(model, _FusedExactGP):
            raise UnsupportedError("MaxPosteriorSampling on the fused path needs a botorch_b200 exact GP model")
        self.model = model
        self.replacement = replacement
        self.chunk_size = int(chunk_size)
        self.joint_max = int(joint_max)  # largest N sampled jointly (N x N covariance: 2 GiB at 16384)
        self.pathwise, self.num_features = bool(pathwise), int(num_features)

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
        if self.pathwise:  # joint draws as sample paths (Matheron's rule): any N
            if observation_noise:
                raise UnsupportedError("pathwise draws are draws of the latent function (no observation noise)")
            paths = MatheronPath(self.model, torch.Size([num_samples]), self.num_features, self.chunk_size)
            return self.maximize_samples(X, paths(X).unsqueeze(-1), num_samples)
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
