"""Hyperparameter fitting on the fused path: exact MLL value + gradient from ``bo_mll_value_grad``.

Mirrors the reference's fitting contract for the default ``SingleTaskGP`` module set:
  * ``ExactMarginalLogLikelihood(likelihood, model)``  gpytorch.mlls (reference call sites: fit.py:75-114,
                                                        optim/closures/model_closures.py:171-184)
  * ``fit_gpytorch_mll``                               botorch/fit.py:75-259 (``_fit_fallback``: attempts, warning policy,
                                                        ``pick_best_of_all_attempts``, ``ModelFittingError``)
  * ``fit_gpytorch_mll_scipy`` / ``scipy_minimize``    botorch/optim/fit.py:47-110, botorch/optim/core.py:55-140
  * ``get_loss_closure_with_grads``                    botorch/optim/closures/model_closures.py:65-104
  * ``get_parameters_and_bounds``                      botorch/optim/utils/model_utils.py:69-109
  * ``sample_all_priors``                              botorch/optim/utils/model_utils.py:153-200
  * priors / constraints                               botorch/models/utils/gpytorch_modules.py:74-127

The O(n^3) arithmetic (Cholesky, inverse factor, W = alpha alpha^T - Khat^-1, the kernel-derivative contraction) runs in
libbotorch_b200; the host adds the prior terms of the d + 2 scalars, divides by n and drives SciPy's L-BFGS-B.
There is no CPU path for the likelihood.
"""
from __future__ import annotations

import math
import logging
import re
import warnings
from dataclasses import dataclass
from enum import Enum, auto
from time import monotonic
from typing import Any, Callable, Dict, Optional, Sequence, Tuple, Union

import numpy as np
import torch
from torch import Tensor
from torch.nn import Module

from . import engine
from .exceptions import BotorchError, NanError, NotPSDError, OptimizationWarning, UnsupportedError
from .models import MIN_INFERRED_NOISE_LEVEL, MIN_LENGTHSCALE, SingleTaskGP

LOG_2PI = math.log(2.0 * math.pi)
MIN_OUTPUTSCALE = 1e-6
# optim/core.py:34-36; SciPy >= 1.15 prints the same messages in upper case ("TOTAL NO. OF ...")
_LBFGSB_MAXITER_MAXFUN_REGEX = re.compile("TOTAL NO. of (ITERATIONS REACHED LIMIT|f AND g EVALUATIONS EXCEEDS LIMIT)", re.IGNORECASE)


class ModelFittingError(BotorchError):
    """botorch/exceptions/errors.py ModelFittingError."""


class OptimizationStatus(int, Enum):  # optim/core.py:39-44
    RUNNING = auto()
    SUCCESS = auto()
    FAILURE = auto()
    STOPPED = auto()


@dataclass
class OptimizationResult:  # optim/core.py:47-52
    step: int
    fval: Union[float, int]
    status: OptimizationStatus
    runtime: Optional[float] = None
    message: Optional[str] = None


# ---------------------------------------------------------------------------------------------------------------------
# priors of the default module set (value and derivative in closed form; d + 2 scalars, host-side plumbing)
# ---------------------------------------------------------------------------------------------------------------------
def _lognormal(x: Tensor, loc: float, scale: float) -> Tuple[Tensor, Tensor]:
    """LogNormal(loc, scale).log_prob(x) and its derivative (gpytorch LogNormalPrior)."""
    lx = x.log()
    lp = -lx - math.log(scale) - 0.5 * LOG_2PI - (lx - loc) ** 2 / (2.0 * scale * scale)
    return lp, -(1.0 + (lx - loc) / (scale * scale)) / x


def _gamma(x: Tensor, concentration: float, rate: float) -> Tuple[Tensor, Tensor]:
    """Gamma(concentration, rate).log_prob(x) and its derivative (gpytorch GammaPrior)."""
    lp = concentration * math.log(rate) + (concentration - 1.0) * x.log() - rate * x - math.lgamma(concentration)
    return lp, (concentration - 1.0) / x - rate


def lengthscale_prior_params(d: int) -> Tuple[float, float]:
    """gpytorch_modules.py:118: LogNormalPrior(loc=sqrt2 + log(d)/2, scale=sqrt3)."""
    return math.sqrt(2.0) + 0.5 * math.log(d), math.sqrt(3.0)


NOISE_PRIOR = (-4.0, 1.0)  # gpytorch_modules.py:88
OUTPUTSCALE_PRIOR = (2.0, 0.15)  # gpytorch_modules.py:49 (GammaPrior of the Scale-Matern factory)


# ---------------------------------------------------------------------------------------------------------------------
# the marginal log likelihood object
# ---------------------------------------------------------------------------------------------------------------------
class ExactMarginalLogLikelihood(Module):
    """``gpytorch.mlls.ExactMarginalLogLikelihood(likelihood, model)`` for the fused ``SingleTaskGP``.

    ``forward()`` returns (log N(y | m, K + D) + sum of prior log-probs) / n as a device scalar; ``value_and_grads()``
    adds the gradient per named parameter.  Parameter names are the reference's (``mll.named_parameters()``:
    ``likelihood.noise_covar.raw_noise``, ``model.mean_module.raw_constant``, ``model.covar_module.raw_lengthscale``).
    """

    def __init__(self, likelihood: Optional[Module], model: SingleTaskGP) -> None:
        super().__init__()
        if not isinstance(model, SingleTaskGP):
            raise UnsupportedError("the fused MLL covers SingleTaskGP (ARD RBF / Matern-5/2, constant mean) only")
        if likelihood is not None and likelihood is not model.likelihood:
            raise UnsupportedError("pass `model.likelihood` (custom likelihoods are outside the fused path)")
        self.likelihood = model.likelihood
        self.model = model
        self._state: Optional[engine.MLLState] = None

    # -- parameters ------------------------------------------------------------------------------------------------
    def _parts(self) -> Dict[str, Optional[Tensor]]:
        m = self.model
        return {
            "lengthscale": m.covar_module.raw_lengthscale,
            "outputscale": getattr(m.covar_module, "raw_outputscale", None),
            "noise": None if m.train_Yvar is not None else m.likelihood.noise_covar.raw_noise,
            "mean": m.mean_module.raw_constant,
        }

    def _mll_state(self) -> engine.MLLState:
        m = self.model
        X = m.train_inputs[0]
        if self._state is None or self._state.gp.train_X.data_ptr() != X.data_ptr():
            self._state = engine.MLLState(X.detach(), m.train_targets.detach(), m._kernel,
                                          with_outputscale=self._parts()["outputscale"] is not None,
                                          noise_vec=m.train_Yvar)  # fmt: skip
        return self._state

    @property
    def evaluations(self) -> int:
        return 0 if self._state is None else self._state.evaluations

    def value_and_grads(self) -> Tuple[Tensor, Dict[str, Tensor]]:
        """(mll, {part: d mll / d part}) with mll = (log marginal + log priors) / n; everything on the device."""
        p = self._parts()
        d = p["lengthscale"].numel()
        n = self.model.train_targets.numel()
        ls = p["lengthscale"].detach().reshape(d)
        value, grad = self._mll_state().evaluate(ls, None if p["noise"] is None else p["noise"].detach(),
                                                 p["mean"].detach(), None if p["outputscale"] is None else p["outputscale"].detach())  # fmt: skip
        total = value[0].clone()
        lp, dlp = _lognormal(ls, *lengthscale_prior_params(d))
        total = total + lp.sum()
        grads = {"lengthscale": ((grad[0, :d] + dlp) / n).reshape(p["lengthscale"].shape)}
        if p["noise"] is not None:
            nz = p["noise"].detach().reshape(())
            lp, dlp = _lognormal(nz, *NOISE_PRIOR)
            total = total + lp
            grads["noise"] = ((grad[0, d + 1] + dlp) / n).reshape(p["noise"].shape)
        if p["outputscale"] is not None:
            os = p["outputscale"].detach().reshape(())
            lp, dlp = _gamma(os, *OUTPUTSCALE_PRIOR)
            total = total + lp
            grads["outputscale"] = ((grad[0, d] + dlp) / n).reshape(p["outputscale"].shape)
        grads["mean"] = (grad[0, d + 2] / n).reshape(p["mean"].shape)
        return total / n, grads

    def forward(self, *args: Any, **kwargs: Any) -> Tensor:
        return self.value_and_grads()[0]


_PART_OF = {
    "likelihood.noise_covar.raw_noise": "noise",
    "model.mean_module.raw_constant": "mean",
    "model.covar_module.raw_lengthscale": "lengthscale",
    "model.covar_module.raw_outputscale": "outputscale",
}


class SumMarginalLogLikelihood(Module):
    """Sum of the exact MLLs of the independent outcomes of a ``ModelListGP`` (gpytorch ``SumMarginalLogLikelihood`` as the
    reference uses it, fit.py:262-285): one ``ExactMarginalLogLikelihood`` per outcome in ``mlls``."""

    def __init__(self, likelihood: Optional[Module], model) -> None:
        super().__init__()
        from .constrained import ModelListGP

        if not isinstance(model, ModelListGP):
            raise UnsupportedError("SumMarginalLogLikelihood takes a botorch_b200 ModelListGP")
        self.model, self.likelihood = model, likelihood
        self.mlls = torch.nn.ModuleList([ExactMarginalLogLikelihood(m.likelihood, m) for m in model.models])

    def forward(self, *args: Any, **kwargs: Any) -> Tensor:
        return sum(mll.forward() for mll in self.mlls)


def get_parameters_and_bounds(mll: ExactMarginalLogLikelihood, requires_grad: Optional[bool] = None):
    """Named parameters and their box bounds (``GreaterThan`` constraints with ``transform=None``: raw == value)."""
    params: Dict[str, Tensor] = {}
    bounds: Dict[str, Tuple[Optional[float], Optional[float]]] = {}
    lower = {"lengthscale": MIN_LENGTHSCALE, "noise": MIN_INFERRED_NOISE_LEVEL, "outputscale": MIN_OUTPUTSCALE}
    parts = mll._parts()
    for name, part in _PART_OF.items():
        p = parts[part]
        if p is None or (requires_grad is not None and p.requires_grad != requires_grad):
            continue
        params[name] = p
        if part in lower:
            bounds[name] = (lower[part], float("inf"))
    return params, bounds


def get_parameters(mll: ExactMarginalLogLikelihood, requires_grad: Optional[bool] = None) -> Dict[str, Tensor]:
    return get_parameters_and_bounds(mll, requires_grad)[0]


def get_loss_closure_with_grads(mll: ExactMarginalLogLikelihood, parameters: Dict[str, Tensor]):
    """Closure () -> (loss = -mll, gradients in the order of ``parameters``); also fills ``param.grad``."""

    def closure(**kwargs: Any) -> Tuple[Tensor, Tuple[Optional[Tensor], ...]]:
        value, grads = mll.value_and_grads()
        out = []
        for name, p in parameters.items():
            g = -grads[_PART_OF[name]].to(p.dtype)
            p.grad = g
            out.append(g)
        return -value, tuple(out)

    return closure


# ---------------------------------------------------------------------------------------------------------------------
# SciPy driver (optim/core.py:55-140, optim/closures/core.py:77-185)
# ---------------------------------------------------------------------------------------------------------------------
class NdarrayOptimizationClosure:
    """ndarray-typed, stateful view of a (value, gradients) closure over named tensors."""

    def __init__(self, closure: Callable[[], Tuple[Tensor, Sequence[Optional[Tensor]]]], parameters: Dict[str, Tensor],
                 fill_value: float = 0.0) -> None:  # fmt: skip
        self.closure = closure
        self.parameters = parameters
        self.fill_value = fill_value
        self._sizes = [p.numel() for p in parameters.values()]

    @property
    def state(self) -> np.ndarray:
        with torch.no_grad():
            flat = torch.cat([p.detach().reshape(-1).to(torch.float64) for p in self.parameters.values()])
        return flat.cpu().numpy()

    @state.setter
    def state(self, x: np.ndarray) -> None:
        dev = next(iter(self.parameters.values())).device
        flat = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64)).to(dev)
        with torch.no_grad():
            for p, chunk in zip(self.parameters.values(), flat.split(self._sizes)):
                p.copy_(chunk.reshape(p.shape).to(p.dtype))

    def __call__(self, state: Optional[np.ndarray] = None, **kwargs: Any) -> Tuple[np.ndarray, np.ndarray]:
        if state is not None:
            self.state = state
        try:
            value, grads = self.closure(**kwargs)
            parts = [value.reshape(1).to(torch.float64)]
            for p, g in zip(self.parameters.values(), grads):
                parts.append(torch.full((p.numel(),), self.fill_value, dtype=torch.float64, device=value.device)
                             if g is None else g.reshape(-1).to(torch.float64))  # fmt: skip
            host = torch.cat(parts).cpu().numpy()  # one device->host read per evaluation
            return host[0], host[1:].copy()
        except (NanError, NotPSDError):  # optim/closures/core.py _handle_numerical_errors: NaN value and gradient
            size = sum(self._sizes)
            return np.full((), np.nan), np.full(size, np.nan)


def scipy_minimize(closure, parameters: Dict[str, Tensor], bounds=None, callback=None, x0: Optional[np.ndarray] = None,
                   method: str = "L-BFGS-B", options: Optional[Dict[str, Any]] = None,
                   timeout_sec: Optional[float] = None) -> OptimizationResult:  # fmt: skip
    from scipy.optimize import minimize

    start_time = monotonic()
    wrapped = closure if isinstance(closure, NdarrayOptimizationClosure) else NdarrayOptimizationClosure(closure, parameters)
    bounds_np = None
    if bounds is not None:
        rows = []
        for name, p in parameters.items():
            lo, hi = bounds.get(name, (None, None))
            lo = -np.inf if lo is None else lo
            hi = np.inf if hi is None else hi
            rows += [(lo, hi)] * p.numel()
        bounds_np = np.asarray(rows, dtype=np.float64)
    steps = [0]

    def wrapped_callback(x: np.ndarray):
        steps[0] += 1
        if timeout_sec is not None and monotonic() - start_time > timeout_sec:
            raise StopIteration
        if callback is not None:
            res = OptimizationResult(step=steps[0], fval=float(wrapped(x)[0]), status=OptimizationStatus.RUNNING,
                                     runtime=monotonic() - start_time)  # fmt: skip
            return callback(parameters, res)

    timed_out = False
    x_init = wrapped.state if x0 is None else np.asarray(x0, dtype=np.float64)
    try:
        raw = minimize(wrapped, x_init, jac=True, bounds=bounds_np, method=method, options=options, callback=wrapped_callback)
        x_fin, fun, nit, success = np.asarray(raw.x), raw.fun, raw.nit, raw.success
        msg = raw.message if isinstance(raw.message, str) else raw.message.decode("ascii")
    except StopIteration:
        timed_out = True
        x_fin, nit, success = wrapped.state, steps[0], False
        fun = float(wrapped(x_fin)[0])
        msg = f"Optimization timed out after {monotonic() - start_time} seconds."
    wrapped.state = x_fin
    if success:
        status = OptimizationStatus.SUCCESS
    else:
        status = (OptimizationStatus.STOPPED if timed_out or _LBFGSB_MAXITER_MAXFUN_REGEX.search(msg)
                  else OptimizationStatus.FAILURE)  # fmt: skip
    return OptimizationResult(fval=fun, step=nit, status=status, message=msg, runtime=monotonic() - start_time)


def fit_gpytorch_mll_scipy(mll: ExactMarginalLogLikelihood, parameters: Optional[Dict[str, Tensor]] = None, bounds=None,
                           closure=None, closure_kwargs: Optional[Dict[str, Any]] = None, method: str = "L-BFGS-B",
                           options: Optional[Dict[str, Any]] = None, callback=None,
                           timeout_sec: Optional[float] = None) -> OptimizationResult:  # fmt: skip
    """optim/fit.py:47-110."""
    _parameters, _bounds = get_parameters_and_bounds(mll)
    bounds = _bounds if bounds is None else {**_bounds, **bounds}
    if parameters is None:
        parameters = {n: p for n, p in _parameters.items() if p.requires_grad}
    if closure is None:
        closure = get_loss_closure_with_grads(mll, parameters)
    if closure_kwargs is not None:
        from functools import partial

        closure = partial(closure, **closure_kwargs)
    result = scipy_minimize(closure=closure, parameters=parameters, bounds=bounds, method=method, options=options,
                            callback=callback, timeout_sec=timeout_sec)  # fmt: skip
    if result.status != OptimizationStatus.SUCCESS:
        warnings.warn(
            f"`scipy_minimize` terminated with status {result.status}, displaying original message from "
            f"`scipy.optimize.minimize`: {result.message}",
            OptimizationWarning,
        )
    return result


# ---------------------------------------------------------------------------------------------------------------------
# fit_gpytorch_mll (fit.py:75-259)
# ---------------------------------------------------------------------------------------------------------------------
def sample_all_priors(model: SingleTaskGP, max_retries: int = 100) -> None:
    """Resample the hyperparameters that carry a prior, rejecting draws outside their constraints
    (optim/utils/model_utils.py:153-200).  torch's global CPU generator, like the reference."""
    d = model.covar_module.raw_lengthscale.numel()

    def draw(param: Tensor, sampler: Callable[[torch.Size], Tensor], lower: float) -> None:
        for i in range(max_retries):
            v = sampler(param.shape)
            if bool((v >= lower).all()):
                with torch.no_grad():
                    param.copy_(v.to(param))
                return
        raise RuntimeError(f"Failed to sample a feasible parameter value from the prior after {max_retries} attempts.")

    loc, scale = lengthscale_prior_params(d)
    draw(model.covar_module.raw_lengthscale, lambda s: torch.exp(loc + scale * torch.randn(s, dtype=torch.float64)), MIN_LENGTHSCALE)
    if model.train_Yvar is None:
        draw(model.likelihood.noise_covar.raw_noise,
             lambda s: torch.exp(NOISE_PRIOR[0] + NOISE_PRIOR[1] * torch.randn(s, dtype=torch.float64)), MIN_INFERRED_NOISE_LEVEL)  # fmt: skip
    os = getattr(model.covar_module, "raw_outputscale", None)
    if os is not None:
        g = torch.distributions.Gamma(torch.tensor(OUTPUTSCALE_PRIOR[0], dtype=torch.float64), OUTPUTSCALE_PRIOR[1])
        draw(os, lambda s: g.sample(s), MIN_OUTPUTSCALE)


def _default_warning_handler(w: warnings.WarningMessage) -> bool:
    """fit.py:51-73 DEFAULT_WARNING_HANDLER over optim/utils/common.py:35-60: True = resolved (no retry).

    * the L-BFGS-B "iteration / evaluation limit reached" message is resolved at DEBUG level: a fit that merely hit
      ``maxiter`` is returned as is, not retried (``_debug_warn``);
    * anything that is not an ``OptimizationWarning`` (e.g. a ``NumericalWarning`` from a jitter retry) and the
      optimiser's time-out message are re-emitted to the caller and resolved (``_rethrow_warn``);
    * every other ``OptimizationWarning`` is unresolved: the attempt failed and the fit is retried from prior samples."""
    message = str(w.message)
    if _LBFGSB_MAXITER_MAXFUN_REGEX.search(message):
        logging.getLogger("botorch").debug(message)
        return True
    if not issubclass(w.category, OptimizationWarning) or "Optimization timed out after" in message:
        warnings.warn_explicit(message, w.category, w.filename, w.lineno)
        return True
    return False


def fit_gpytorch_mll(mll: ExactMarginalLogLikelihood, closure=None, optimizer: Optional[Callable] = None,
                     closure_kwargs: Optional[Dict[str, Any]] = None, optimizer_kwargs: Optional[Dict[str, Any]] = None,
                     max_attempts: int = 5, pick_best_of_all_attempts: bool = False,
                     warning_handler: Callable[[warnings.WarningMessage], bool] = _default_warning_handler,
                     caught_exception_types: Tuple[type, ...] = (NotPSDError,), **ignore: Any) -> ExactMarginalLogLikelihood:  # fmt: skip
    """Maximise the exact MLL over the hyperparameters with L-BFGS-B; retry from prior samples on optimiser warnings.

    Returns ``mll`` in eval mode on success; raises ``ModelFittingError`` when every attempt failed.  A
    ``SumMarginalLogLikelihood`` over a ``ModelListGP`` is fitted one outcome at a time (``_fit_list``, fit.py:262-285)."""
    import copy

    if isinstance(mll, SumMarginalLogLikelihood):
        mll.train()
        for sub_mll in mll.mlls:
            fit_gpytorch_mll(sub_mll, closure=closure, optimizer=optimizer, closure_kwargs=closure_kwargs,
                             optimizer_kwargs=optimizer_kwargs, max_attempts=max_attempts,
                             pick_best_of_all_attempts=pick_best_of_all_attempts, warning_handler=warning_handler,
                             caught_exception_types=caught_exception_types, **ignore)  # fmt: skip
        return mll.eval() if not any(sub_mll.training for sub_mll in mll.mlls) else mll

    optimizer = fit_gpytorch_mll_scipy if optimizer is None else optimizer
    optimizer_kwargs = {} if optimizer_kwargs is None else optimizer_kwargs
    mll.train()
    if closure is None:
        closure = get_loss_closure_with_grads(mll, get_parameters(mll, requires_grad=True))
    if closure_kwargs is not None:
        from functools import partial

        closure = partial(closure, **closure_kwargs)
    start = copy.deepcopy({k: v.detach().clone() for k, v in mll.model.state_dict().items()})
    best_mll, best_state = -float("inf"), None
    for attempt in range(1, 1 + max_attempts):
        if attempt > 1:
            mll.model.load_state_dict(start)  # roll back, then resample the free parameters
            mll.train()
            sample_all_priors(mll.model)
        try:
            with warnings.catch_warnings(record=True) as warning_list:
                warnings.simplefilter("always", category=OptimizationWarning)
                result = optimizer(mll, closure=closure, **optimizer_kwargs)
            success = True
            for w in warning_list:
                if not warning_handler(w):
                    warnings.warn_explicit(str(w.message), w.category, w.filename, w.lineno)
                    success = False
            if success and not pick_best_of_all_attempts:
                mll.eval()
                return mll
            if success and -result.fval > best_mll:
                best_mll = -result.fval
                best_state = {k: v.detach().clone() for k, v in mll.model.state_dict().items()}
        except caught_exception_types:
            pass
    if best_state is not None:
        mll.model.load_state_dict(best_state)
        mll.eval()
        return mll
    mll.model.load_state_dict(start)
    mll.train()
    raise ModelFittingError("All attempts to fit the model have failed.")
