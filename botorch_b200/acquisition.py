"""qLogNoisyExpectedImprovement on the fused CUDA path.

Contract mirrored from the reference:
  * ``AcquisitionFunction``            botorch/acquisition/acquisition.py:25-74
  * ``qLogNoisyExpectedImprovement``   botorch/acquisition/logei.py:237-501 (constructor arguments, buffers
                                        ``X_baseline`` / ``baseline_samples`` / ``_baseline_best_f``, attributes
                                        ``_baseline_L`` / ``sampler`` / ``base_sampler`` / ``q_in`` / ``tau_relu`` / ``tau_max``)
  * forward decorators                 botorch/utils/transforms.py:228-336 (t-batch handling, pending points, ensemble
                                        logmeanexp, output-shape check)
  * base-sample bookkeeping            botorch/acquisition/cached_cholesky.py:167-186, botorch/sampling/normal.py:68-131
  * failure conventions                botorch/acquisition/cached_cholesky.py:139-153 (warning, never an exception)

``forward`` is backed by a ``torch.autograd.Function`` whose forward and backward are the two C-ABI calls
``bo_qlognei_fwd`` / ``bo_qlognei_bwd``; ``torch.autograd.grad`` (generation/gen.py:211) therefore runs the
hand-written reverse kernels.  Nothing here computes on the CPU.
"""
from __future__ import annotations

import math
import warnings
from copy import deepcopy
from typing import Callable, Optional

import torch
from torch import Tensor
from torch.nn import Module

from . import engine
from .exceptions import BotorchWarning, NanError, NotPSDError, NumericalWarning, NumericsWarning, UnsupportedError
from .posteriors import MCMC_DIM
from .models import Model, _FusedExactGP
from .sampling import MCSampler, get_sampler

TAU_RELU = 1e-6  # logei.py:66
TAU_MAX = 1e-2  # logei.py:67
DEFAULT_MC_SAMPLES = 512  # acquisition/acquisition.py:116


class AcquisitionFunction(Module):
    """Abstract acquisition function (acquisition.py:25-74)."""

    _log: bool = False

    def __init__(self, model: Model) -> None:
        super().__init__()
        self.model: Model = model

    def set_X_pending(self, X_pending: Optional[Tensor] = None) -> None:
        if X_pending is not None:
            if X_pending.requires_grad:
                warnings.warn(
                    "Pending points require a gradient but the acquisition function will not provide a gradient to "
                    "these points.",
                    BotorchWarning,
                    stacklevel=2,
                )
            self.X_pending = X_pending.detach().clone()
        else:
            self.X_pending = None

    def forward(self, X: Tensor) -> Tensor:  # pragma: no cover - abstract
        raise NotImplementedError


class IdentityMCObjective(Module):
    """objective(samples) = samples[..., 0] for a single-outcome model (acquisition/objective.py:282-298)."""

    def forward(self, samples: Tensor, X: Optional[Tensor] = None) -> Tensor:
        return samples.squeeze(-1)


class LinearMCObjective(Module):
    """objective(samples) = sum_i weights[i] samples[..., i] (acquisition/objective.py:301-341).  The fused path takes
    the single-outcome case: the weight is applied inside the kernels (``bo_nei.obj_weight``), e.g. ``weights=[-1.]``
    turns the acquisition functions into their minimisation counterparts."""

    def __init__(self, weights: Tensor) -> None:
        super().__init__()
        if weights.dim() != 1:
            raise ValueError("weights must be a one-dimensional tensor.")
        self.register_buffer("weights", weights)

    def forward(self, samples: Tensor, X: Optional[Tensor] = None) -> Tensor:
        if samples.shape[-1] != self.weights.shape[-1]:
            raise RuntimeError("Output shape of samples not equal to that of weights")
        return torch.einsum("...m, m", [samples, self.weights])


class GenericMCObjective(Module):
    """Objective from a callable ``f(samples, X)`` mapping ``sample_shape x batch_shape x q x m`` samples (and the
    ``batch_shape x q x d`` inputs) to ``sample_shape x batch_shape x q`` objective values (acquisition/objective.py:344-384).
    Arbitrary Python cannot run inside the fused kernels: acquisition functions given one evaluate it with torch on the
    materialised samples, between the sample kernels and the reduction kernels (``botorch_b200/constrained.py``)."""

    def __init__(self, objective: Callable[[Tensor, Optional[Tensor]], Tensor]) -> None:
        super().__init__()
        self.objective = objective

    def forward(self, samples: Tensor, X: Optional[Tensor] = None) -> Tensor:
        return self.objective(samples, X=X)


def _needs_sample_path(model, objective, constraints) -> bool:
    """True when the call cannot use the all-in-kernel path of a single model: outcome constraints or a generic objective
    need the samples in HBM and torch between the kernels (the model-list classes of ``constrained.py``)."""
    from .constrained import ModelListGP

    return isinstance(model, ModelListGP) or constraints is not None or isinstance(objective, GenericMCObjective)


def _as_model_list(model):
    from .constrained import ModelListGP

    if isinstance(model, ModelListGP):
        return model
    if not isinstance(model, _FusedExactGP) or model._is_fully_bayesian:
        raise UnsupportedError("outcome constraints / generic objectives on the fused path need exact (non-ensemble) "
                               "botorch_b200 GP models")  # fmt: skip
    return ModelListGP(model)


def _objective_weight(objective) -> float:
    """The scalar the kernels multiply the samples with; anything but identity / one-outcome linear is off the path."""
    if objective is None or isinstance(objective, IdentityMCObjective):
        return 1.0
    if isinstance(objective, LinearMCObjective) and objective.weights.numel() == 1:
        w = float(objective.weights[0])
        if w == 0.0 or not math.isfinite(w):
            raise ValueError("the objective weight must be finite and non-zero")
        return w
    raise UnsupportedError(
        f"objective {type(objective).__name__} is not on the fused path (SURVEY.md section 8f): use None, "
        "IdentityMCObjective or a one-outcome LinearMCObjective"
    )


def legacy_ei_numerics_warning(legacy_name: str) -> None:
    """NumericsWarning pointing from a legacy EI class to its LogEI counterpart (exceptions/warnings.py:67-99)."""
    logei_name = {"qExpectedImprovement": "qLogExpectedImprovement",
                  "qNoisyExpectedImprovement": "qLogNoisyExpectedImprovement"}.get(legacy_name)  # fmt: skip
    if logei_name is not None:
        warnings.warn(
            f"{legacy_name} has known numerical issues that lead to suboptimal optimization performance. It is "
            f"strongly recommended to simply replace\n\n\t {legacy_name} \t --> \t {logei_name} \n\n"
            "instead, which fixes the issues and has the same API. See https://arxiv.org/abs/2310.20708 for details.",
            NumericsWarning,
            stacklevel=3,
        )


def prune_inferior_points(
    model: Model,
    X: Tensor,
    objective=None,
    posterior_transform=None,
    constraints=None,
    num_samples: int = 2048,
    max_frac: float = 1.0,
    sampler: Optional[MCSampler] = None,
    marginalize_dim: Optional[int] = None,
) -> Tensor:
    """Keep the points of ``X`` (n x d) that are the best one in at least one joint posterior sample
    (acquisition/utils.py:245-349).  The joint draw over the n points runs on the GPU (blocked Cholesky + GEMM)."""
    for name, arg in (("posterior_transform", posterior_transform), ("constraints", constraints)):
        if arg is not None:
            raise UnsupportedError(f"`{name}` is not on the fused path yet (SURVEY.md section 8f): pass None")
    obj_w = _objective_weight(objective)
    if marginalize_dim is None and model._is_ensemble:
        marginalize_dim = MCMC_DIM
    if X.ndim > 2:
        raise UnsupportedError("Batched inputs `X` are currently unsupported by prune_inferior_points")
    if X.size(-2) == 0:
        raise ValueError("X must have at least one point.")
    if max_frac <= 0 or max_frac > 1.0:
        raise ValueError(f"max_frac must take values in (0, 1], is {max_frac}")
    max_points = math.ceil(max_frac * X.size(-2))
    with torch.no_grad():
        posterior = model.posterior(X=X)
        if sampler is None:
            sampler = get_sampler(posterior=posterior, sample_shape=torch.Size([num_samples]))
        samples = sampler(posterior)
    obj_vals = obj_w * samples.squeeze(-1)  # IdentityMCObjective / one-outcome LinearMCObjective
    if obj_vals.ndim > 2:
        if obj_vals.ndim == 3 and marginalize_dim is not None:
            if marginalize_dim < 0:  # MCMC_DIM counts from the end of the model's batch shape + (q, m)
                marginalize_dim = 1 + (marginalize_dim % obj_vals.ndim)
            obj_vals = obj_vals.mean(dim=marginalize_dim)
        else:
            raise UnsupportedError("Models with multiple batch dims are currently unsupported by prune_inferior_points.")
    is_best = torch.argmax(obj_vals, dim=-1)
    idcs, counts = torch.unique(is_best, return_counts=True)
    if len(idcs) > max_points:  # (sic) the reference indexes by the sort order of the unique list here
        counts, order_idcs = torch.sort(counts, descending=True)
        idcs = order_idcs[:max_points]
    return X[idcs]


class _ShapeOnlyPosterior:
    """What a sampler needs to size its base samples for r jointly sampled points (no arithmetic)."""

    batch_range = (0, -1)

    def __init__(self, model_batch: torch.Size, r: int, device, dtype) -> None:
        self.base_sample_shape = torch.Size(model_batch) + torch.Size([r])
        self.batch_shape = torch.Size(model_batch)
        self.device, self.dtype = device, dtype


class _QLogNEIFunction(torch.autograd.Function):
    """acq[M x B] = fused qLogNEI; backward = the hand-derived reverse kernels."""

    @staticmethod
    def forward(ctx, X: Tensor, nei: engine.NEIState, Zq: Tensor, tau_relu: float, tau_max: float):
        with_grad = ctx.needs_input_grad[0]
        Xd = X.detach().to(torch.float64).contiguous()
        B, q, d = Xd.shape
        ge = nei.graphed(B, q, d, Zq, tau_relu, tau_max, with_grad)
        if ge is not None and not ge.busy:  # replay the captured launch sequence
            acq, info = ge.run_forward(Xd)
            ctx.ge = ge if with_grad else None
        else:
            acq, info, work = nei.forward(Xd, Zq, tau_relu, tau_max, with_grad=with_grad)
            ctx.ge = None
            if with_grad:
                ctx.nei, ctx.work, ctx.taus = nei, work, (tau_relu, tau_max)
                ctx.save_for_backward(Xd, Zq)
        ctx.in_dtype = X.dtype
        ctx.mark_non_differentiable(info)
        return acq.to(X.dtype), info

    @staticmethod
    def backward(ctx, grad_acq: Tensor, _grad_info):
        if ctx.ge is not None:
            g = ctx.ge.run_backward(grad_acq.to(torch.float64))
            return g.to(ctx.in_dtype), None, None, None, None
        Xd, Zq = ctx.saved_tensors
        g = ctx.nei.backward(Xd, Zq, ctx.taus[0], ctx.taus[1], grad_acq.to(torch.float64).contiguous(), ctx.work)
        ctx.work = None  # the workspace may be recycled from here on
        return g.sum(dim=0).to(ctx.in_dtype), None, None, None, None


class _FusedImprovementBase(AcquisitionFunction):
    """Shared forward of the improvement-based MC acquisition functions (SampleReducingMCAcquisitionFunction.forward,
    monte_carlo.py:253-289): subclasses provide ``self._nei`` (device state), ``self._set_sampler(q)`` (-> ``self._Zq``)
    and the class attribute ``_log``."""

    tau_relu, tau_max = TAU_RELU, TAU_MAX
    _info_host, _info_event, _info_pending = None, None, False

    def _check_fused_args(self, model, **unsupported) -> None:
        if not isinstance(model, _FusedExactGP):
            raise UnsupportedError(
                f"the fused {type(self).__name__} needs a botorch_b200 exact GP model, got {type(model).__name__}; "
                "no generic fallback"
            )
        for name, arg in unsupported.items():
            if arg is not None:
                raise UnsupportedError(f"`{name}` is not on the fused path yet (SURVEY.md section 8f): pass None")

    def _reduction_mode(self) -> int:
        if not self._log:
            return engine.REDUCE_PLAIN
        return engine.REDUCE_LOG if getattr(self, "_fat", True) else engine.REDUCE_LOG_SOFT

    # -- forward --------------------------------------------------------------------------------------
    def forward(self, X: Tensor) -> Tensor:
        """X: (b) x q x d  ->  (b) acquisition values (one per t-batch)."""
        if X.dim() < 2:
            raise ValueError(
                f"{type(self).__name__} requires X to have at least 2 dimensions, but received X with only "
                f"{X.dim()} dimensions."
            )
        if self.X_pending is not None:  # concatenate_pending_points
            pend = self.X_pending.expand(*X.shape[:-2], *self.X_pending.shape[-2:]).to(X)
            X = torch.cat([X, pend], dim=-2)
        X = X if X.dim() > 2 else X.unsqueeze(0)  # add the t-batch dimension
        batch_shape = X.shape[:-2]
        q, d = X.shape[-2:]
        Xf = X.reshape(-1, q, d)
        if Xf.device != self._nei.device:
            raise engine._lib.BoError(
                f"X is on {Xf.device}; the acquisition function lives on {self._nei.device}. There is no CPU path."
            )
        self._set_sampler(q_in=q)
        acq, info = _QLogNEIFunction.apply(Xf, self._nei, self._Zq, self.tau_relu, self.tau_max)  # taus unused when not _log
        if self._strict_numerics or not (torch.is_grad_enabled() and Xf.requires_grad):
            # no gradient is being recorded (raw-sample sweeps, final evaluations: the caller reads the values next anyway),
            # or a caller asked for it: look at the failure flags of THIS call now and resolve them before returning
            failed = self._failed_tbatches(info)
            if failed is not None and not self._nei.c.no_baseline:
                acq = self._forward_resolved(Xf, failed, acq)
            elif failed is not None:  # fixed-incumbent family: no cached baseline root to fall back from
                warnings.warn("The q x q posterior covariance could not be factored; the affected t-batches were "
                              "evaluated without their root.", BotorchWarning, stacklevel=3)  # fmt: skip
        else:
            self._warn_on_failures(info)
        if self.model._is_ensemble:  # t_batch_mode_transform: (log)mean over the hyperparameter samples
            acq = torch.logsumexp(acq, dim=0) - math.log(acq.shape[0]) if self._log else acq.mean(dim=0)
        else:
            acq = acq[0]
        return acq.reshape(batch_shape)

    # -- numerical failures: the reference's fallback to standard sampling (cached_cholesky.py:139-165) ---------------------
    _strict_numerics = False
    last_call_failed = False

    def strict_numerics(self, on: bool = True) -> None:
        """Check the failure flags of every call before it returns, also while a gradient is recorded (one host wait per
        forward).  ``optim.HostEvaluator`` switches it on to redo an evaluation whose deferred check failed."""
        self._strict_numerics = bool(on)

    def _failed_tbatches(self, info: Tensor) -> Optional[Tensor]:
        """Indices of the t-batches whose cached update failed in the call that produced ``info`` (waits for it), or None.
        Jitter retries that succeeded (info > 0) only warn, as psd_safe_cholesky does."""
        if info.numel() == 2 and info.dim() == 1:  # graph replay: (min, max) of the flags; the full array is static
            lo, hi = (int(v) for v in info.tolist())
            full = None
        else:
            lo, hi = int(info.min()), int(info.max())
            full = info
        if hi > 0:
            warnings.warn(
                f"A not p.d., added jitter of {1e-8 * 10 ** (hi - 1):.1e} to the diagonal", NumericalWarning, stacklevel=4
            )
        if lo >= 0:
            return None
        if full is None:
            full = self._nei.last_graph_info
        return torch.nonzero((full < 0).any(dim=0)).view(-1)

    def _forward_resolved(self, Xf: Tensor, failed: Tensor, acq: Tensor) -> Tensor:
        """Re-evaluate the t-batches ``failed`` the way the reference does when ``sample_cached_cholesky`` raises
        ``NanError`` / ``NotPSDError``: warn, then sample from the JOINT posterior over [X_baseline; X]
        (``get_posterior_samples`` -> ``psd_safe_cholesky`` of the (n_b+q) covariance) and keep the last q rows.

        The joint factor with jitter delta on its diagonal is [[L_b', 0], [Sigma_qb L_b'^-T, chol(Schur' + delta I)]] with
        L_b' = chol(Sigma_bb + delta I): the cached update with the baseline block re-factored under delta and the q x q
        block under the same delta.  So each delta of psd_safe_cholesky's schedule (0, 1e-8 ... 1e-3) gets a temporary
        baseline state and the SAME fused kernels (values and gradients); ``best_f`` stays the cached one, as in the
        reference (``_baseline_best_f`` is not recomputed on this path).  A t-batch that fails every delta raises
        ``NotPSDError`` (``NanError`` for NaN inputs) like ``psd_safe_cholesky``.  The reference falls back for the whole
        call; the t-batches that did not fail are identical on both routes (delta = 0), so only the failed ones move."""
        warnings.warn(
            "Low-rank cholesky updates failed due NaNs or due to an ill-conditioned covariance matrix. "
            "Falling back to standard sampling.",
            BotorchWarning,
            stacklevel=4,
        )
        if bool(torch.isnan(Xf[failed]).any()):
            raise NanError(f"cholesky: {int(torch.isnan(Xf[failed]).sum())} of {Xf[failed].numel()} elements are NaN.")
        nei = self._nei
        for ge in nei.__dict__.get("_graphs", {}).values():
            ge.busy = False  # the failed forward is dropped below: its backward will never run
        # the call that failed leaves the autograd graph altogether (its backward would push NaNs through the broken state):
        # the t-batches that were fine are evaluated again on their own
        B = Xf.shape[0]
        ok_mask = torch.ones(B, dtype=torch.bool, device=Xf.device)
        ok_mask[failed] = False
        ok = torch.nonzero(ok_mask).view(-1)
        out = torch.zeros_like(acq.detach())
        if ok.numel():
            a_ok, _ = _QLogNEIFunction.apply(Xf[ok].contiguous(), nei, self._Zq, self.tau_relu, self.tau_max)
            out = out.index_copy(1, ok, a_ok)
        todo = failed
        for i in range(-1, engine.CHOLESKY_MAX_TRIES):
            delta = 0.0 if i < 0 else engine.CHOLESKY_JITTER * (10**i)
            if i >= 0:
                warnings.warn(f"A not p.d., added jitter of {delta:.1e} to the diagonal", NumericalWarning, stacklevel=4)
            tmp = engine.NEIState(nei.gp, nei.X_b, nei.Z_b, reduction=nei.c.reduction, obj_weight=nei.c.obj_weight,
                                  forced_jitter=delta)  # fmt: skip
            if tmp.setup_failed:
                continue
            tmp.best_f_pad.copy_(nei.best_f_pad)
            a, info = _QLogNEIFunction.apply(Xf[todo].contiguous(), tmp, self._Zq, self.tau_relu, self.tau_max)
            if info.numel() != a.numel():  # graph replay hands over (min, max) only: read the static array
                info = tmp.last_graph_info
            good = (info.reshape(a.shape) >= 0).all(dim=0)
            idx = todo[good]
            if idx.numel():
                out = out.index_copy(1, idx, a[:, good])
            todo = todo[~good]
            if todo.numel() == 0:
                return out
        raise NotPSDError(f"Matrix not positive definite after repeatedly adding jitter up to {delta:.1e}.")

    def _warn_on_failures(self, info: Tensor) -> None:
        """Queue the failure flags of this call; report the flags of the previous one.

        Reading ``info`` right away would stall the host on the forward kernels before it can enqueue the loss and the
        backward (one idle gap per L-BFGS-B evaluation).  The (min, max) pair goes to pinned host memory behind an
        event instead and is examined at the next call -- by then the caller has read the loss, so the event is
        long complete -- or by ``check_numerics()``, which ``optimize_acqf`` calls before returning.
        """
        self.check_numerics()
        if self._info_host is None:
            self._info_host = torch.empty(2, dtype=torch.int32).pin_memory()
            self._info_event = torch.cuda.Event()
        if info.numel() != 2 or info.dim() != 1:  # eager path: per-(model, t-batch) flags; graphs hand over (min, max)
            info = torch.stack(torch.aminmax(info)).to(torch.int32)
        self._info_host.copy_(info, non_blocking=True)
        self._info_event.record()
        self._info_pending = True

    def check_numerics(self, quiet: bool = False) -> None:
        """Emit the warnings of the most recent evaluation (waits for it if it is still running); ``last_call_failed`` tells
        whether its cached update failed (``quiet``: only set the flag)."""
        if not self._info_pending:
            return
        self._info_event.synchronize()
        self._info_pending = False
        lo, hi = int(self._info_host[0]), int(self._info_host[1])
        self.last_call_failed = lo < 0
        if quiet:
            return
        if hi > 0:
            warnings.warn(
                f"A not p.d., added jitter of {1e-8 * 10 ** (hi - 1):.1e} to the diagonal", NumericalWarning, stacklevel=3
            )
        if lo < 0:
            warnings.warn(
                "Low-rank cholesky updates failed due NaNs or due to an ill-conditioned covariance matrix in a call "
                "whose flags were read late (a gradient was being recorded); its affected t-batches were evaluated "
                "without their q x q root.  gen_candidates_scipy re-evaluates such a call through the reference's "
                "fallback; use acqf.strict_numerics() to check every call before it returns.",
                BotorchWarning,
                stacklevel=3,
            )


class qLogNoisyExpectedImprovement(_FusedImprovementBase):
    """MC batch log noisy expected improvement, evaluated by one fused GPU pipeline per call.

    qLogNEI(X) ~ log E[ max(max_j f(x_j) - max f(X_baseline), 0) ] with the smooth fat-tailed approximations of
    [Ament2023logei]: log_fatplus(tau_relu) for the ReLU, fatmax(tau_max) over the q points, logmeanexp over samples.
    """

    _log = True

    def __new__(cls, model=None, *args, **kwargs):
        if cls is qLogNoisyExpectedImprovement:  # several outcomes / outcome constraints / generic objective: constrained.py
            from .constrained import qLogNEIModelList

            objective = kwargs.get("objective", args[2] if len(args) > 2 else None)
            constraints = kwargs.get("constraints", args[5] if len(args) > 5 else None)
            if _needs_sample_path(model, objective, constraints):
                return qLogNEIModelList(_as_model_list(model), *args, **kwargs)
        return super().__new__(cls)

    def __init__(
        self,
        model: Model,
        X_baseline: Tensor,
        sampler: Optional[MCSampler] = None,
        objective=None,
        posterior_transform=None,
        X_pending: Optional[Tensor] = None,
        constraints=None,
        eta=1e-3,
        fat: bool = True,
        prune_baseline: bool = False,
        cache_root: bool = True,
        tau_max: float = TAU_MAX,
        tau_relu: float = TAU_RELU,
        marginalize_dim: Optional[int] = None,
        *,
        _factor_setup: bool = True,
    ) -> None:
        super().__init__(model=model)
        if not isinstance(model, _FusedExactGP):
            raise UnsupportedError(
                f"the fused qLogNEI needs a botorch_b200 exact GP model, got {type(model).__name__}; no generic fallback"
            )
        for name, arg in (("posterior_transform", posterior_transform), ("constraints", constraints)):
            if arg is not None:
                raise UnsupportedError(f"`{name}` is not on the fused path yet (SURVEY.md section 8f): pass None")
        self._obj_weight = _objective_weight(objective)
        self.objective = objective if objective is not None else IdentityMCObjective()
        if X_baseline.dim() != 2:
            raise UnsupportedError("X_baseline must be r x d (one baseline set shared by all t-batches)")
        self.tau_relu, self.tau_max = float(tau_relu), float(tau_max)
        self._fat = bool(fat)  # False: log_softplus / smooth_amax instead of log_fatplus / fatmax (logei.py:122,532)
        # cache_root=False only changes the order of operations in the reference (joint Cholesky of the
        # (n_b+q) block); the block Cholesky computed here is that same factor, so both settings share the path.
        self._cache_root = bool(cache_root)
        self.sample_shape = torch.Size([DEFAULT_MC_SAMPLES]) if sampler is None else sampler.sample_shape
        self.set_X_pending(X_pending)
        if prune_baseline:  # logei.py:377-385
            X_baseline = prune_inferior_points(model=model, X=X_baseline, objective=objective,
                                               marginalize_dim=marginalize_dim)
        self._init_baseline(X_baseline, sampler, _factor_setup)

    # -- one-time setup (logei.py:364-420) ------------------------------------------------------------
    def _init_baseline(self, X_baseline: Tensor, sampler: Optional[MCSampler], setup: bool = True) -> None:
        gp: engine.GPState = self.model.gp_state
        self.register_buffer("X_baseline", X_baseline.detach())
        n_b = X_baseline.shape[-2]
        shape_post = _ShapeOnlyPosterior(self.model.batch_shape, n_b, gp.device, X_baseline.dtype)
        if sampler is None:
            sampler = get_sampler(shape_post, sample_shape=self.sample_shape)
        if len(sampler.sample_shape) != 1:
            raise UnsupportedError("the fused path takes a one-dimensional sample_shape")
        self.sampler = sampler
        S = self.sample_shape[0]

        def draw_base_samples() -> Tensor:  # host: a scrambled Sobol engine of dimension n_b (S x (1 per model-batch dim) x n_b)
            self.sampler._construct_base_samples(shape_post)
            return self.sampler.base_samples.reshape(S, n_b)

        # setup=False: allocate only; botorch_b200.distributed broadcasts the factors from the rank that computed them.
        # Otherwise the device factors the baseline posterior while the host draws the base samples (engine.NEIState).
        self._nei = engine.NEIState(gp, self.X_baseline.to(gp.device), draw_base_samples if setup else draw_base_samples(),
                                    setup=setup, reduction=self._reduction_mode(), obj_weight=self._obj_weight, S=S)
        self.base_sampler = deepcopy(self.sampler)
        self.q_in = -1
        self._Zq: Optional[Tensor] = None
        self.register_buffer("_baseline_best_f", None)
        self._refresh_baseline_views()

    def _refresh_baseline_views(self) -> None:
        """(Re)derive the reference-named views of the device state (after setup or after a factor broadcast)."""
        best = self._nei.best_f  # [M x S]
        best = best.mT if self.model._is_ensemble else best[0]
        self._baseline_best_f = best.to(self.X_baseline.dtype).clone()
        L = self._nei.baseline_L
        self._baseline_L = (L if self.model._is_ensemble else L[0]).to(self.X_baseline.dtype)

    @property
    def baseline_samples(self) -> Tensor:
        """S x (M) x n_b x 1 posterior samples at X_baseline (materialised on request only)."""
        nei = self._nei
        Z = nei.Z_b.unsqueeze(1).expand(nei.S, nei.gp.M, nei.n_b)
        out = engine.root_apply(nei.baseline_L.contiguous(), nei.mean_b[:, : nei.n_b], Z)
        out = out if self.model._is_ensemble else out[:, 0]
        return out.unsqueeze(-1).to(self.X_baseline.dtype)

    @property
    def baseline_obj(self) -> Tensor:
        return self._obj_weight * self.baseline_samples.squeeze(-1)

    def compute_best_f(self, obj: Tensor) -> Tensor:
        val = self._baseline_best_f
        view = torch.Size([val.shape[0], *(1,) * (obj.ndim - val.ndim - 1), *val.shape[1:]])
        return val.view(view).to(obj)

    # -- base samples of the q new points (cached_cholesky.py:167-186) -----------------------------------
    def _set_sampler(self, q_in: int) -> None:
        if self.q_in == q_in and self._Zq is not None:
            return
        n_b = self.X_baseline.shape[-2]
        gp = self._nei.gp
        # the t-batch dimension of X_full = [X_baseline; X] is collapsed to 1 by the sampler
        if getattr(self, "_prefetched_q", None) != q_in:
            post = _ShapeOnlyPosterior(torch.Size([1]) + self.model.batch_shape, n_b + q_in, gp.device, self.X_baseline.dtype)
            self.sampler._update_base_samples(posterior=post, base_sampler=self.base_sampler)
        self._prefetched_q = None
        S = self.sample_shape[0]
        Z = self.sampler.base_samples.reshape(S, n_b + q_in)
        self._Zq = Z[:, n_b:].to(device=gp.device, dtype=torch.float64).contiguous()
        self.q_in = q_in


    def prefetch_base_samples(self, q_in: int):
        """Start drawing the base samples of ``q_in`` new points on a host thread and return it (``join`` it before the first
        forward).  A scrambled ``SobolEngine`` of dimension n_b + q takes tens of milliseconds to construct on the host
        (45 ms at n_b = 2000) and is independent of everything else ``optimize_acqf`` does before its first evaluation --
        e.g. drawing the raw samples, another Sobol engine.  Host work only; the device copy stays in ``_set_sampler``."""
        import threading

        if self.q_in == q_in and self._Zq is not None:
            return None
        n_b = self.X_baseline.shape[-2]
        post = _ShapeOnlyPosterior(torch.Size([1]) + self.model.batch_shape, n_b + q_in, self._nei.gp.device,
                                   self.X_baseline.dtype)  # fmt: skip

        def work() -> None:
            self.sampler._update_base_samples(posterior=post, base_sampler=self.base_sampler)
            self._prefetched_q = q_in

        t = threading.Thread(target=work, daemon=True)
        t.start()
        return t


class qNoisyExpectedImprovement(qLogNoisyExpectedImprovement):
    """MC batch noisy expected improvement, qNEI(X) = E[max(max_j f(x_j) - max f(X_baseline), 0)]
    (acquisition/monte_carlo.py:417-620): the sample path of qLogNEI with ``clamp_min(0)``, ``amax`` over q and the
    plain mean over samples.  ``prune_baseline`` defaults to True as in the reference."""

    _log = False

    def __new__(cls, model=None, *args, **kwargs):
        if cls is qNoisyExpectedImprovement:
            from .constrained import qNEIModelList

            objective = kwargs.get("objective", args[2] if len(args) > 2 else None)
            constraints = kwargs.get("constraints", args[7] if len(args) > 7 else None)
            if _needs_sample_path(model, objective, constraints):
                return qNEIModelList(_as_model_list(model), *args, **kwargs)
        return object.__new__(cls)

    def __init__(
        self,
        model: Model,
        X_baseline: Tensor,
        sampler: Optional[MCSampler] = None,
        objective=None,
        posterior_transform=None,
        X_pending: Optional[Tensor] = None,
        prune_baseline: bool = True,
        cache_root: bool = True,
        constraints=None,
        eta=1e-3,
        marginalize_dim: Optional[int] = None,
        *,
        _factor_setup: bool = True,
    ) -> None:
        legacy_ei_numerics_warning(type(self).__name__)
        super().__init__(model, X_baseline, sampler=sampler, objective=objective, posterior_transform=posterior_transform,
                         X_pending=X_pending, constraints=constraints, eta=eta, prune_baseline=prune_baseline,
                         cache_root=cache_root, marginalize_dim=marginalize_dim, _factor_setup=_factor_setup)  # fmt: skip


class qLogExpectedImprovement(_FusedImprovementBase):
    """MC batch log expected improvement over a fixed incumbent ``best_f`` (acquisition/logei.py:138-234):
    joint posterior over the q points, log_fatplus(f - best_f), fatmax over q, logmeanexp over samples."""

    _log = True

    def __new__(cls, model=None, *args, **kwargs):
        if cls in (qLogExpectedImprovement, qExpectedImprovement):  # constraints / generic objective: constrained.py
            from .constrained import qEIModelList, qLogEIModelList

            objective = kwargs.get("objective", args[2] if len(args) > 2 else None)
            constraints = kwargs.get("constraints", args[5] if len(args) > 5 else None)
            if _needs_sample_path(model, objective, constraints):
                target = qLogEIModelList if cls is qLogExpectedImprovement else qEIModelList
                return target(_as_model_list(model), *args, **kwargs)
        return object.__new__(cls)

    def __init__(
        self,
        model: Model,
        best_f,
        sampler: Optional[MCSampler] = None,
        objective=None,
        posterior_transform=None,
        X_pending: Optional[Tensor] = None,
        constraints=None,
        eta=1e-3,
        fat: bool = True,
        tau_max: float = TAU_MAX,
        tau_relu: float = TAU_RELU,
    ) -> None:
        AcquisitionFunction.__init__(self, model=model)
        self._check_fused_args(model, posterior_transform=posterior_transform, constraints=constraints)
        self._fat = bool(fat)
        self.tau_relu, self.tau_max = float(tau_relu), float(tau_max)
        self._obj_weight = _objective_weight(objective)
        self.objective = objective if objective is not None else IdentityMCObjective()
        self._init_fixed_incumbent(best_f, sampler, X_pending)

    def _init_fixed_incumbent(self, best_f, sampler: Optional[MCSampler], X_pending: Optional[Tensor]) -> None:
        gp: engine.GPState = self.model.gp_state
        self.register_buffer("best_f", torch.as_tensor(best_f, dtype=torch.float64))
        self.sample_shape = torch.Size([DEFAULT_MC_SAMPLES]) if sampler is None else sampler.sample_shape
        if len(self.sample_shape) != 1:
            raise UnsupportedError("the fused path takes a one-dimensional sample_shape")
        self.sampler = sampler
        self.set_X_pending(X_pending)
        self._nei = engine.EIState(gp, self.best_f.to(gp.device), self.sample_shape[0],
                                   reduction=self._reduction_mode(), obj_weight=self._obj_weight)  # fmt: skip
        self.q_in, self._Zq = -1, None

    def _set_sampler(self, q_in: int) -> None:
        """Base samples S x 1 x q of the joint draw, shared by all t-batches and models (sampling/base.py:78-125)."""
        if self.q_in == q_in and self._Zq is not None:
            return
        gp = self._nei.gp
        post = _ShapeOnlyPosterior(torch.Size([1]) + self.model.batch_shape, q_in, gp.device, torch.float64)
        if self.sampler is None:  # MCAcquisitionFunction.get_posterior_samples, monte_carlo.py:116-125
            self.sampler = get_sampler(post, sample_shape=self.sample_shape)
        self.sampler._construct_base_samples(post)
        S = self.sample_shape[0]
        self._Zq = self.sampler.base_samples.reshape(S, q_in).to(device=gp.device, dtype=torch.float64).contiguous()
        self.q_in = q_in


class qExpectedImprovement(qLogExpectedImprovement):
    """MC batch expected improvement, qEI(X) = E[max(max_j f(x_j) - best_f, 0)] (acquisition/monte_carlo.py:333-414)."""

    _log = False

    def __init__(self, model: Model, best_f, sampler: Optional[MCSampler] = None, objective=None,
                 posterior_transform=None, X_pending: Optional[Tensor] = None, constraints=None, eta=1e-3) -> None:  # fmt: skip
        legacy_ei_numerics_warning(type(self).__name__)
        AcquisitionFunction.__init__(self, model=model)
        self._check_fused_args(model, posterior_transform=posterior_transform, constraints=constraints)
        self._obj_weight = _objective_weight(objective)
        self.objective = objective if objective is not None else IdentityMCObjective()
        self._init_fixed_incumbent(best_f, sampler, X_pending)
