"""Outcome constraints on the fused path: several single-output exact GPs (``ModelListGP``) under one qLogNEI
(SURVEY.md 8f N2).

Mirrors
  * ``ModelListGP``                                   botorch/models/model_list_gp_regression.py (independent single-output models)
  * ``SampleReducingMCAcquisitionFunction`` + constraints   acquisition/monte_carlo.py:253-330 (``_apply_constraints``: log improvement +
                                                      log feasibility before the q-reduction)
  * ``compute_smoothed_feasibility_indicator``        utils/objective.py:134-180 (``log_fatmoid`` / ``logexpit`` of -constraint/eta)
  * ``compute_best_feasible_objective``               acquisition/utils.py:90-163 (incumbent over the FEASIBLE baseline samples;
                                                      ``_estimate_objective_lower_bound`` when a sample has none)
  * base samples of a multi-output posterior          sampling/normal.py:34-131, utils/low_rank.py:43-82 (S x 1 x (n_b+q) x m)

Every output runs the same fused sample pipeline as the single-model qLogNEI (``bo_qlognei_fwd`` / ``bo_qlognei_bwd``);
what couples the outputs -- the constraint callables, arbitrary Python on an S x B x q x m sample tensor -- is evaluated
by torch on the device between ``bo_mc_reduce_fwd`` / ``bo_mc_reduce_bwd``, which hold the reductions.
"""
from __future__ import annotations

import math
from typing import Callable, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor
from torch.quasirandom import SobolEngine

from . import _lib, engine
from ._lib import check, pad8, ptr
from .acquisition import (DEFAULT_MC_SAMPLES, TAU_MAX, TAU_RELU, AcquisitionFunction, GenericMCObjective, IdentityMCObjective,
                          LinearMCObjective, _FusedImprovementBase)  # fmt: skip
from .exceptions import UnsupportedError
from .models import Model, _FusedExactGP
from .sampling import (IIDNormalSampler, MCSampler, SobolQMCNormalSampler, _normal_from_uniform, draw_sobol_normal_samples,
                       manual_seed, sobol_draw)  # fmt: skip

_INV_SQRT_3 = math.sqrt(1.0 / 3.0)  # utils/safe_math.py: inflection point of the fatmoid


def fatmoid(X: Tensor, tau: Union[float, Tensor] = 1.0) -> Tensor:
    """utils/safe_math.py:429-446: C^2 Heaviside approximation with an O(1/x^2) tail."""
    X = X / tau
    cauchy = lambda x: 1 / (1 + x.square())
    return torch.where(X < 0, 2 / 3 * cauchy(X - _INV_SQRT_3), 1 - 2 / 3 * cauchy(X + _INV_SQRT_3))


def log_fatmoid(X: Tensor, tau: Union[float, Tensor] = 1.0) -> Tensor:
    return fatmoid(X, tau=tau).log()


def logexpit(X: Tensor) -> Tensor:
    """utils/safe_math.py:97-99: log sigmoid = -log1pexp(-X) (log1pexp: softplus with its linear branch)."""
    return -torch.nn.functional.softplus(-X, threshold=math.log(torch.finfo(X.dtype).max) / 2)


def compute_feasibility_indicator(constraints, samples: Tensor) -> Tensor:
    ind = torch.ones(samples.shape[:-1], dtype=torch.bool, device=samples.device)
    for constraint in constraints or []:
        ind = ind.logical_and(constraint(samples) <= 0)
    return ind


def compute_smoothed_feasibility_indicator(constraints, samples: Tensor, eta, log: bool = False, fat: bool = False) -> Tensor:
    if type(eta) is not Tensor:
        eta = torch.full((len(constraints),), eta)
    if len(eta) != len(constraints):
        raise ValueError("Number of provided constraints and number of provided etas do not match.")
    if not (eta > 0).all():
        raise ValueError("eta must be positive.")
    is_feasible = torch.zeros_like(samples[..., 0])
    log_sigmoid = log_fatmoid if fat else logexpit
    for constraint, e in zip(constraints, eta):
        is_feasible = is_feasible + log_sigmoid(-constraint(samples) / float(e))
    return is_feasible if log else is_feasible.exp()


class _ListPosterior:
    """mean / variance of independent outputs side by side (``... x q x m``)."""

    def __init__(self, posts) -> None:
        self.mean = torch.cat([p.mean for p in posts], dim=-1)
        self.variance = torch.cat([p.variance for p in posts], dim=-1)
        self.device, self.dtype = self.mean.device, self.mean.dtype


class ModelListGP(Model):
    """Independent single-output fused GPs modelling m outcomes of the same inputs."""

    def __init__(self, *gp_models: _FusedExactGP) -> None:
        super().__init__()
        for m in gp_models:
            if not isinstance(m, _FusedExactGP) or m._is_fully_bayesian:
                raise UnsupportedError("ModelListGP on the fused path takes single-output botorch_b200 exact GPs")
        self.models = torch.nn.ModuleList(gp_models)
        self.likelihood = torch.nn.ModuleList([m.likelihood for m in gp_models])  # LikelihoodList of the reference

    @classmethod
    def from_batched(cls, train_X: Tensor, train_Y: Tensor, train_Yvar: Optional[Tensor] = None, likelihood=None,
                     covar_module=None, mean_module=None, outcome_transform="DEFAULT", input_transform=None, *,
                     lengthscale=None, noise=None, mean_constant=0.0, outputscale=None, kernel: str = "rbf") -> "ModelListGP":  # fmt: skip
        """What ``SingleTaskGP(train_X, train_Y)`` means for ``train_Y`` of shape n x m, m > 1: the reference turns the
        outputs into a batch of m independent GPs over the same inputs with per-output hyperparameters and per-output
        standardisation (models/gpytorch.py:405-466, models/gp_regression.py:164-254, Standardize(m=m)) -- exactly m
        single-output models.  Per-output hyperparameters: ``lengthscale`` m x d (or d, shared), ``noise`` /
        ``mean_constant`` / ``outputscale`` of length m (or scalars)."""
        from .models import SingleTaskGP

        m = train_Y.shape[-1]
        if not (outcome_transform is None or isinstance(outcome_transform, str)):
            raise UnsupportedError("pass outcome_transform='DEFAULT' (per-output Standardize) or None for multi-output data")

        def col(v, k, per_output_dim):
            if v is None or isinstance(v, (int, float)):
                return v
            t = torch.as_tensor(v)
            return t[k] if t.dim() > per_output_dim else t

        return cls(*[
            SingleTaskGP(train_X, train_Y[:, k : k + 1], None if train_Yvar is None else train_Yvar[:, k : k + 1], likelihood,
                         covar_module, mean_module, outcome_transform, input_transform, lengthscale=col(lengthscale, k, 1),
                         noise=col(noise, k, 0), mean_constant=col(mean_constant, k, 0), outputscale=col(outputscale, k, 0),
                         kernel=kernel)
            for k in range(m)
        ])  # fmt: skip

    @property
    def num_outputs(self) -> int:
        return len(self.models)

    @property
    def batch_shape(self) -> torch.Size:
        return torch.Size()

    def posterior(self, X: Tensor, output_indices: Optional[List[int]] = None, observation_noise=False,
                  posterior_transform=None) -> _ListPosterior:  # fmt: skip
        if posterior_transform is not None:
            raise UnsupportedError("posterior_transform is outside the fused path")
        idx = range(len(self.models)) if output_indices is None else output_indices
        return _ListPosterior([self.models[i].posterior(X, observation_noise=observation_noise) for i in idx])


def _one_outcome(objective, m: int) -> Tuple[int, float]:
    """(index, weight) of an objective that reads exactly one outcome linearly -- the form the reduction kernels apply
    themselves.  A ``GenericMCObjective`` returns (-1, 1.0): its values are computed by torch on the sample tensor."""
    if isinstance(objective, GenericMCObjective):
        return -1, 1.0
    if m == 1 and (objective is None or isinstance(objective, IdentityMCObjective)):
        return 0, 1.0
    if not isinstance(objective, LinearMCObjective):
        raise UnsupportedError(
            "with several outcomes the fused path needs `objective=LinearMCObjective(weights)` selecting ONE outcome "
            "(one nonzero weight) or a `GenericMCObjective`; constraints may read every outcome"
        )
    w = objective.weights.detach().reshape(-1).cpu().double()
    nz = torch.nonzero(w).reshape(-1)
    if w.numel() != m or nz.numel() != 1:
        raise UnsupportedError(f"objective weights must have length {m} with exactly one nonzero entry, got {w.tolist()}")
    return int(nz[0]), float(w[nz[0]])


def _objective_values(objective, obj_idx: int, obj_w: float, samples: Tensor, X: Optional[Tensor]) -> Tensor:
    """objective(samples, X) for ``... x m`` samples: the linear one-outcome form, or the generic callable."""
    if obj_idx >= 0:
        return obj_w * samples[..., obj_idx]
    return objective(samples, X=X)


def prune_inferior_points_list(model: ModelListGP, X: Tensor, objective, constraints=None, num_samples: int = 2048,
                               max_frac: float = 1.0, sampler: Optional[MCSampler] = None) -> Tensor:  # fmt: skip
    """``prune_inferior_points`` (acquisition/utils.py:245-349) over several independent outcomes: keep the points of ``X``
    that are the best FEASIBLE one in at least one joint posterior sample.  One joint draw over the (points x outcomes)
    block as the reference's sampler makes it; per outcome the n x n posterior covariance, its blocked Cholesky and the
    root-times-base-samples GEMM run on the GPU; infeasible entries are set below the worst objective (:336-340)."""
    if X.ndim > 2:
        raise UnsupportedError("Batched inputs `X` are currently unsupported by prune_inferior_points")
    if X.size(-2) == 0:
        raise ValueError("X must have at least one point.")
    if max_frac <= 0 or max_frac > 1.0:
        raise ValueError(f"max_frac must take values in (0, 1], is {max_frac}")
    m, n = model.num_outputs, X.size(-2)
    obj_idx, obj_w = _one_outcome(objective, m)
    max_points = math.ceil(max_frac * n)
    if sampler is None:
        cls = SobolQMCNormalSampler if n * m <= SobolEngine.MAXDIM else IIDNormalSampler
        sampler = cls(sample_shape=torch.Size([num_samples]))
    S = int(sampler.sample_shape[0])
    if isinstance(sampler, SobolQMCNormalSampler):
        z = draw_sobol_normal_samples(d=n * m, n=S, dtype=torch.float64, seed=sampler.seed)
    else:
        with manual_seed(seed=sampler.seed):
            z = torch.randn(S, n * m, dtype=torch.float64)
    z = z.view(S, n, m).to(X.device)
    with torch.no_grad():
        cols = []
        for k, mod in enumerate(model.models):
            post = mod.posterior(X)
            base = z[..., k].reshape(torch.Size([S]) + post.base_sample_shape)
            cols.append(post.rsample_from_base_samples(torch.Size([S]), base).reshape(S, n))
        samples = torch.stack(cols, dim=-1)  # S x n x m
        obj_vals = _objective_values(objective, obj_idx, obj_w, samples, X).clone()
        infeas = ~compute_feasibility_indicator(constraints, samples)
        if bool(infeas.any()):
            obj_vals[infeas] = obj_vals.min() - 1
        is_best = torch.argmax(obj_vals, dim=-1)
    idcs, counts = torch.unique(is_best, return_counts=True)
    if len(idcs) > max_points:  # (sic) the reference indexes by the sort order of the unique list here
        counts, order_idcs = torch.sort(counts, descending=True)
        idcs = order_idcs[:max_points]
    return X[idcs]


class _ConstrainedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, X: Tensor, acqf: "qLogNEIModelList"):
        Xd = X.detach().to(torch.float64).contiguous()
        B, q, _ = Xd.shape
        N, S = B * q, acqf.S
        Zq = acqf._base_samples_q(q)
        need_grad = bool(ctx.needs_input_grad[0])  # no-grad calls (raw samples) skip the backward's workspace regions
        works, views, infos = [], [], []
        for k, nei in enumerate(acqf._neis):  # 1. cached-Cholesky samples of every output (fused pipeline, reductions ignored)
            _, info, work = nei.forward(Xd, Zq[k], acqf.tau_relu, acqf.tau_max, with_grad=need_grad)
            works.append(work)
            infos.append(info.reshape(-1))
            views.append(nei.samples_view(work, B, q))
        # jitter / failed q x q roots of any outcome are reported like the fused path does: (min, max) of the flags, deferred
        acqf._warn_on_failures(torch.stack(torch.aminmax(torch.cat(infos))).to(torch.int32))
        samples = torch.stack([v[:, :N] for v in views], dim=-1).reshape(S, B, q, len(views))
        ldn = pad8(N)
        lw = None
        if acqf._constraints is not None:  # 2. log feasibility weights (arbitrary callables: torch on the device)
            lw = torch.zeros(S, ldn, dtype=torch.float64, device=Xd.device)
            with torch.no_grad():
                lw[:, :N] = compute_smoothed_feasibility_indicator(acqf._constraints, samples, acqf._eta, log=True,
                                                                   fat=acqf._fat).reshape(S, N)  # fmt: skip
        acq = torch.empty(B, dtype=torch.float64, device=Xd.device)
        lse = torch.empty(B, dtype=torch.float64, device=Xd.device)
        if acqf._obj_idx >= 0:  # linear one-outcome objective: the kernel reads that outcome's samples in place
            Fobj = views[acqf._obj_idx]
        else:  # generic objective (acquisition/objective.py:344-384): arbitrary callable, torch on the device
            Fobj = torch.zeros(S, ldn, dtype=torch.float64, device=Xd.device)
            with torch.no_grad():
                Fobj[:, :N] = acqf.objective(samples, X=Xd).reshape(S, N)
        lib = _lib.lib()
        check(lib.bo_mc_reduce_fwd(acqf._ctx, S, B, q, Fobj.stride(0), ptr_view(Fobj), ptr(lw), ptr(acqf._best_f),
                                   acqf._obj_w, acqf.tau_relu, acqf.tau_max, acqf._reduction, ptr(acq), ptr(lse),
                                   _lib.stream_ptr(Xd.device)))  # fmt: skip
        ctx.acqf, ctx.works, ctx.views, ctx.Zq = acqf, works, views, Zq
        ctx.Fobj = Fobj
        ctx.lw, ctx.lse, ctx.in_dtype = lw, lse, X.dtype
        ctx.save_for_backward(Xd, samples)
        return acq.to(X.dtype)

    @staticmethod
    def backward(ctx, grad_acq: Tensor):
        acqf: "qLogNEIModelList" = ctx.acqf
        Xd, samples = ctx.saved_tensors
        B, q, _ = Xd.shape
        N, S, m = B * q, acqf.S, len(ctx.views)
        ldn = pad8(N)
        dev = Xd.device
        Fobj = ctx.Fobj
        dF = torch.zeros(S, Fobj.stride(0), dtype=torch.float64, device=dev)
        lw_bar = None if ctx.lw is None else torch.zeros(S, Fobj.stride(0), dtype=torch.float64, device=dev)
        lw_arg = None
        if ctx.lw is not None:  # the reduce kernels index lw with the pitch of the sample rows
            lw_arg = ctx.lw if ldn == Fobj.stride(0) else torch.nn.functional.pad(ctx.lw, (0, Fobj.stride(0) - ldn))
        lib = _lib.lib()
        check(lib.bo_mc_reduce_bwd(acqf._ctx, S, B, q, Fobj.stride(0), ptr_view(Fobj), ptr(lw_arg), ptr(acqf._best_f),
                                   acqf._obj_w, acqf.tau_relu, acqf.tau_max, acqf._reduction, ptr(ctx.lse),
                                   ptr(grad_acq.to(torch.float64).contiguous()), ptr(dF), ptr(lw_bar),
                                   _lib.stream_ptr(dev)))  # fmt: skip
        gs = None
        if ctx.lw is not None:  # chain d loss / d lw through the constraint callables into every output's samples
            with torch.enable_grad():
                s = samples.detach().requires_grad_(True)
                lw = compute_smoothed_feasibility_indicator(acqf._constraints, s, acqf._eta, log=True, fat=acqf._fat)
                if lw.requires_grad:
                    (gs,) = torch.autograd.grad(lw, s, grad_outputs=lw_bar[:, :N].reshape(S, B, q), allow_unused=True)
                else:  # constraints that do not read the samples differentiably (constants, hard thresholds)
                    gs = None
        grad = None
        if acqf._obj_idx < 0:  # chain d loss / d objective through the generic callable into the samples and into X
            with torch.enable_grad():
                s = samples.detach().requires_grad_(True)
                Xg = Xd.detach().requires_grad_(True)
                obj = acqf.objective(s, X=Xg)
                if obj.requires_grad:
                    go, gx = torch.autograd.grad(obj, [s, Xg], grad_outputs=dF[:, :N].reshape(S, B, q), allow_unused=True)
                    if go is not None:
                        gs = go if gs is None else gs + go
                    grad = gx
        for k, nei in enumerate(acqf._neis):
            dFk = torch.zeros(S, N, dtype=torch.float64, device=dev) if gs is None else gs[..., k].reshape(S, N).clone()
            if k == acqf._obj_idx:
                dFk += dF[:, :N]
            if k != acqf._obj_idx and not bool(dFk.any()):
                continue  # this output does not reach the loss
            ctx.views[k][:, :N] = dFk
            gk = nei.backward_external(Xd, ctx.Zq[k], acqf.tau_relu, acqf.tau_max, ctx.works[k])[0]
            grad = gk if grad is None else grad + gk
        ctx.works = ctx.views = ctx.Fobj = None
        if grad is None:
            grad = torch.zeros_like(Xd)
        return grad.to(ctx.in_dtype), None


def ptr_view(t: Tensor) -> int:
    """Device pointer of a row-pitched view into a workspace (rows contiguous, pitch = stride(0))."""
    assert t.is_cuda and t.dtype == torch.float64 and t.stride(-1) == 1
    return t.data_ptr()


class qLogNEIModelList(AcquisitionFunction):
    """qLogNoisyExpectedImprovement over a ``ModelListGP`` with outcome constraints (what
    ``qLogNoisyExpectedImprovement(model_list, X_baseline, objective=..., constraints=[...])`` constructs)."""

    _log = True
    _reduction = engine.REDUCE_LOG
    _info_host, _info_event, _info_pending = None, None, False
    _warn_on_failures = _FusedImprovementBase._warn_on_failures
    check_numerics = _FusedImprovementBase.check_numerics

    def __init__(self, model: ModelListGP, X_baseline: Tensor, sampler: Optional[MCSampler] = None, objective=None,
                 posterior_transform=None, X_pending: Optional[Tensor] = None,
                 constraints: Optional[Sequence[Callable[[Tensor], Tensor]]] = None, eta: Union[Tensor, float] = 1e-3,
                 fat: bool = True, prune_baseline: bool = False, cache_root: bool = True, tau_max: float = TAU_MAX,
                 tau_relu: float = TAU_RELU, marginalize_dim: Optional[int] = None) -> None:  # fmt: skip
        super().__init__(model=model)
        if posterior_transform is not None or marginalize_dim is not None:
            raise UnsupportedError("posterior_transform / marginalize_dim are not on the constrained fused path")
        if X_baseline.dim() != 2:
            raise UnsupportedError("X_baseline must be r x d (one baseline set shared by all t-batches)")
        m = model.num_outputs
        self._obj_idx, self._obj_w = _one_outcome(objective, m)
        if prune_baseline:  # logei.py:377-385
            X_baseline = prune_inferior_points_list(model, X_baseline, objective, constraints=constraints)
        self.objective = objective
        self._constraints = list(constraints) if constraints else None
        self._eta, self._fat = eta, bool(fat)
        if not self._fat and self._reduction == engine.REDUCE_LOG:  # fat = False: log_softplus / smooth_amax (logei.py:122,532)
            self._reduction = engine.REDUCE_LOG_SOFT
        self.tau_relu, self.tau_max = float(tau_relu), float(tau_max)
        self.set_X_pending(X_pending)
        self.register_buffer("X_baseline", X_baseline.detach())
        self.sampler = sampler if sampler is not None else SobolQMCNormalSampler(torch.Size([DEFAULT_MC_SAMPLES]))
        if len(self.sampler.sample_shape) != 1 or not isinstance(self.sampler, (SobolQMCNormalSampler, IIDNormalSampler)):
            raise UnsupportedError("the fused path takes a Sobol or iid normal sampler with a one-dimensional sample_shape")
        self.S = int(self.sampler.sample_shape[0])
        gps = [mod.gp_state for mod in model.models]
        dev = gps[0].device
        self._ctx = gps[0].ctx
        n_b = X_baseline.shape[0]
        self._Zb = self._draw(n_b, m)  # S x n_b x m: one draw over the (points x outcomes) block, like the reference
        Xb = self.X_baseline.to(dev)
        self._neis = [engine.NEIState(gp, Xb, self._Zb[:, :, k].to(dev).contiguous(), reduction=self._reduction,
                                      obj_weight=(self._obj_w if k == self._obj_idx else 1.0)).keep_samples()
                      for k, gp in enumerate(gps)]  # fmt: skip
        self._Zq_cache = {}
        # incumbent: best FEASIBLE objective over the baseline samples of all outcomes (acquisition/utils.py:90-163)
        base = torch.stack([engine.root_apply(nei.baseline_L.contiguous(), nei.mean_b[:, : nei.n_b],
                                              nei.Z_b.unsqueeze(1).expand(nei.S, 1, nei.n_b))[:, 0] for nei in self._neis], dim=-1)  # fmt: skip
        self.baseline_samples = base  # S x n_b x m
        self.baseline_obj = _objective_values(self.objective, self._obj_idx, self._obj_w, base, Xb)
        self._best_f = self._compute_best_feasible_objective(base, self.baseline_obj).to(torch.float64).contiguous()
        self._baseline_best_f = self._best_f

    def _draw(self, rows: int, m: int, first_row: int = 0) -> Tensor:
        """S x rows x m standard-normal base samples, generated as the reference's sampler does for a collapsed
        ``S x 1 x rows x m`` posterior (sampling/normal.py:178-209 Sobol, :133-175 iid).  ``first_row`` > 0: rows before it
        are not needed by the caller and are left at zero (their Sobol dimensions are not computed)."""
        if isinstance(self.sampler, SobolQMCNormalSampler) and first_row > 0:
            cols = torch.arange(first_row * m, rows * m)
            z = torch.zeros(self.S, rows * m, dtype=torch.float64)
            u = sobol_draw(rows * m, self.S, seed=self.sampler.seed, dtype=torch.float64, columns=cols)
            z[:, cols] = _normal_from_uniform(u)
        elif isinstance(self.sampler, SobolQMCNormalSampler):
            z = draw_sobol_normal_samples(d=rows * m, n=self.S, dtype=torch.float64, seed=self.sampler.seed)
        else:
            with manual_seed(seed=self.sampler.seed):
                z = torch.randn(self.S, rows * m, dtype=torch.float64)
        return z.view(self.S, rows, m)

    def _base_samples_q(self, q: int) -> List[Tensor]:
        """Per outcome S x q base samples of the new points: redraw for n_b + q rows, keep the baseline rows fixed
        (``_update_base_samples``, sampling/normal.py:88-131) -- only the last q rows are used."""
        if q not in self._Zq_cache:
            n_b, m = self._Zb.shape[1], self._Zb.shape[2]
            dev = self._neis[0].device
            full = self._draw(n_b + q, m, first_row=n_b)
            self._Zq_cache[q] = [full[:, n_b:, k].to(dev).contiguous() for k in range(m)]
        return self._Zq_cache[q]

    def _compute_best_feasible_objective(self, samples: Tensor, obj: Tensor) -> Tensor:
        if self._constraints is None:
            return obj.amax(dim=-1)
        is_feasible = compute_feasibility_indicator(self._constraints, samples)  # S x n_b
        if bool(is_feasible.any(dim=-1).all()):
            infeasible_value = -float("inf")
        else:  # _estimate_objective_lower_bound: 6-sigma lower bound over 32 convex combinations of the baseline points
            Xb = self.X_baseline.to(samples.device)
            wts = torch.rand(32, Xb.shape[-2], dtype=Xb.dtype, device=Xb.device)
            wts = wts / wts.sum(dim=0, keepdim=True)
            post = self.model.posterior(wts @ Xb)
            lb = _objective_values(self.objective, self._obj_idx, self._obj_w,
                                   post.mean - 6 * post.variance.clamp_min(0).sqrt(), wts @ Xb)
            infeasible_value = float(-(-(lb.min().clamp_max(0.0))))
        return torch.where(is_feasible, obj, torch.full_like(obj, infeasible_value)).amax(dim=-1)

    def forward(self, X: Tensor) -> Tensor:
        if X.dim() < 2:
            raise ValueError(f"{type(self).__name__} requires X to have at least 2 dimensions, got {X.dim()}.")
        if self.X_pending is not None:
            pend = self.X_pending.expand(*X.shape[:-2], *self.X_pending.shape[-2:]).to(X)
            X = torch.cat([X, pend], dim=-2)
        X = X if X.dim() > 2 else X.unsqueeze(0)
        batch_shape, (q, d) = X.shape[:-2], X.shape[-2:]
        Xf = X.reshape(-1, q, d)
        if Xf.device != self._neis[0].device:
            raise _lib.BoError(f"X is on {Xf.device}; the acquisition function lives on {self._neis[0].device}. No CPU path.")
        return _ConstrainedFn.apply(Xf, self).reshape(batch_shape)


class qNEIModelList(qLogNEIModelList):
    """``qNoisyExpectedImprovement(model_list, X_baseline, objective=..., constraints=[...])`` (acquisition/monte_carlo.py:
    417-620 with ``_apply_constraints`` :305-330): relu improvement times the sigmoid feasibility weights, ``amax`` over q,
    plain mean over the samples -- the BO_REDUCE_PLAIN branch of the same reduction kernels.  ``prune_baseline`` defaults to
    True and the weights are plain sigmoids (``fat=False``) as in the reference."""

    _log = False
    _reduction = engine.REDUCE_PLAIN

    def __init__(self, model: ModelListGP, X_baseline: Tensor, sampler: Optional[MCSampler] = None, objective=None,
                 posterior_transform=None, X_pending: Optional[Tensor] = None, prune_baseline: bool = True,
                 cache_root: bool = True, constraints=None, eta: Union[Tensor, float] = 1e-3,
                 marginalize_dim: Optional[int] = None) -> None:  # fmt: skip
        from .acquisition import legacy_ei_numerics_warning

        legacy_ei_numerics_warning("qNoisyExpectedImprovement")
        super().__init__(model, X_baseline, sampler=sampler, objective=objective, posterior_transform=posterior_transform,
                         X_pending=X_pending, constraints=constraints, eta=eta, fat=False, prune_baseline=prune_baseline,
                         cache_root=cache_root, marginalize_dim=marginalize_dim)  # fmt: skip


class qLogEIModelList(qLogNEIModelList):
    """``qLogExpectedImprovement(model, best_f, objective=..., constraints=[...])`` (acquisition/logei.py:138-234 with
    ``_apply_constraints``, monte_carlo.py:305-330) when outcome constraints or a generic objective are present: the joint
    posterior over the q candidate points of every outcome (``engine.EIState``: no baseline block), the fixed incumbent
    ``best_f``, log improvement + log feasibility, the q-reduction and logmeanexp in the same reduction kernels."""

    def __init__(self, model: ModelListGP, best_f, sampler: Optional[MCSampler] = None, objective=None,
                 posterior_transform=None, X_pending: Optional[Tensor] = None,
                 constraints: Optional[Sequence[Callable[[Tensor], Tensor]]] = None, eta: Union[Tensor, float] = 1e-3,
                 fat: bool = True, tau_max: float = TAU_MAX, tau_relu: float = TAU_RELU) -> None:  # fmt: skip
        AcquisitionFunction.__init__(self, model=model)
        if posterior_transform is not None:
            raise UnsupportedError("posterior_transform is not on the constrained fused path")
        m = model.num_outputs
        self._obj_idx, self._obj_w = _one_outcome(objective, m)
        self.objective = objective
        self._constraints = list(constraints) if constraints else None
        self._eta, self._fat = eta, bool(fat)
        if not self._fat and self._reduction == engine.REDUCE_LOG:
            self._reduction = engine.REDUCE_LOG_SOFT
        self.tau_relu, self.tau_max = float(tau_relu), float(tau_max)
        self.set_X_pending(X_pending)
        gps = [mod.gp_state for mod in model.models]
        dev = gps[0].device
        self._ctx = gps[0].ctx
        self.register_buffer("best_f", torch.as_tensor(best_f, dtype=torch.float64))
        if self.best_f.numel() != 1:
            raise UnsupportedError("the fused path takes a scalar best_f for non-batched models")
        self.sampler = sampler
        self.sample_shape = torch.Size([DEFAULT_MC_SAMPLES]) if sampler is None else sampler.sample_shape
        if len(self.sample_shape) != 1:
            raise UnsupportedError("the fused path takes a one-dimensional sample_shape")
        if sampler is not None and not isinstance(sampler, (SobolQMCNormalSampler, IIDNormalSampler)):
            raise UnsupportedError("the fused path takes a Sobol or iid normal sampler")
        self.S = int(self.sample_shape[0])
        self._best_f = self.best_f.reshape(1).expand(self.S).to(dev).contiguous()  # per-sample incumbent, all equal
        self._neis = [engine.EIState(gp, self.best_f.to(dev), self.S, reduction=self._reduction,
                                     obj_weight=(self._obj_w if k == self._obj_idx else 1.0)).keep_samples()
                      for k, gp in enumerate(gps)]  # fmt: skip
        self._Zq_cache = {}

    def _base_samples_q(self, q: int) -> List[Tensor]:
        """Per outcome S x q base samples: ONE draw over the (q points x m outcomes) block of the joint posterior
        (sampling/normal.py:34-86; get_sampler: Sobol when q * m fits the engine, iid otherwise)."""
        if q not in self._Zq_cache:
            m = len(self._neis)
            if self.sampler is None:
                cls = SobolQMCNormalSampler if q * m <= SobolEngine.MAXDIM else IIDNormalSampler
                self.sampler = cls(sample_shape=self.sample_shape)
            dev = self._neis[0].device
            full = self._draw(q, m)
            self._Zq_cache[q] = [full[:, :, k].to(dev).contiguous() for k in range(m)]
        return self._Zq_cache[q]


class qEIModelList(qLogEIModelList):
    """``qExpectedImprovement(model, best_f, objective=..., constraints=[...])`` (acquisition/monte_carlo.py:333-414): relu
    improvement times the plain sigmoid feasibility weights, ``amax`` over q, mean over the samples."""

    _log = False
    _reduction = engine.REDUCE_PLAIN

    def __init__(self, model: ModelListGP, best_f, sampler: Optional[MCSampler] = None, objective=None,
                 posterior_transform=None, X_pending: Optional[Tensor] = None, constraints=None,
                 eta: Union[Tensor, float] = 1e-3) -> None:  # fmt: skip
        from .acquisition import legacy_ei_numerics_warning

        legacy_ei_numerics_warning("qExpectedImprovement")
        super().__init__(model, best_f, sampler=sampler, objective=objective, posterior_transform=posterior_transform,
                         X_pending=X_pending, constraints=constraints, eta=eta, fat=False)
