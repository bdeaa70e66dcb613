"""Multi-start optimisation of an acquisition function (host control flow around the fused kernels).

Contract mirrored from the reference:
  * ``optimize_acqf`` / batch path          botorch/optim/optimize.py:246-394, 397-543
  * ``gen_batch_initial_conditions``        botorch/optim/initializers.py:243-438
  * ``initialize_q_batch``                  botorch/optim/initializers.py:893-965
  * ``gen_candidates_scipy``                botorch/generation/gen.py:46-298 (L-BFGS-B over the flattened restarts,
                                             value + ``torch.autograd.grad`` per function evaluation)
  * ``get_best_candidates``                 botorch/generation/gen.py:427-455
  * ``columnwise_clamp`` / scipy bounds     botorch/optim/utils/common.py, botorch/optim/parameter_constraints.py:53-96

The SciPy L-BFGS-B driver stays on the host exactly as in the reference; every acquisition value and
gradient it consumes comes from the CUDA path.  The raw-sample and restart batches are the sharding seams used by
``botorch_b200.distributed`` (``init_batch_limit`` / ``batch_limit``, optimize.py:268-317).
"""
from __future__ import annotations

import ctypes as C
import math
import time
import warnings
from typing import Callable, Dict, Optional, Tuple, Union

import numpy as np
import torch
from scipy import optimize as _sciopt
from torch import Tensor
from torch.quasirandom import SobolEngine

from .exceptions import (
    BadInitialCandidatesWarning,
    OptimizationWarning,
    UnsupportedError,
)
from .sampling import draw_sobol_samples, manual_seed

INIT_OPTION_KEYS = {
    "alpha", "batch_limit", "eta", "init_batch_limit", "nonnegative", "n_burnin", "sample_around_best",
    "sample_around_best_sigma", "sample_around_best_prob_perturb", "sample_around_best_subset_sigma", "seed",
    "thinning",
}  # fmt: skip


# -------------------------------------------------------------------------------------------------
# initial conditions
# -------------------------------------------------------------------------------------------------
def initialize_q_batch_indices(Y: Tensor, n: int, eta: float = 1.0) -> Tensor:
    """Indices ``initialize_q_batch`` selects: ``n`` of the raw q-batches with probability ~ exp(eta * zscore(Y)), the best
    one always kept (initializers.py:906-951).  Consumes torch's host RNG stream exactly as the reference does."""
    n_samples = Y.shape[0]
    if n > n_samples:
        raise RuntimeError(f"n ({n}) cannot be larger than the number of provided samples ({n_samples})")
    if n == n_samples:
        return torch.arange(n_samples, device=Y.device)
    Ystd = Y.std(dim=0)
    if bool(torch.any(Ystd == 0)):
        warnings.warn(
            "All acquisition values for raw samples points are the same for at least one batch. Choosing initial "
            "conditions at random.",
            BadInitialCandidatesWarning,
        )
        return torch.randperm(n=n_samples, device=Y.device)[:n]
    max_idx = torch.argmax(Y, dim=0)
    etaZ = eta * (Y - Y.mean(dim=0)) / Ystd
    weights = torch.exp(etaZ)
    while bool(torch.isinf(weights).any()):
        etaZ = etaZ * 0.5
        weights = torch.exp(etaZ)
    idcs = torch.multinomial(weights, n)
    if max_idx not in idcs:
        idcs[-1] = max_idx
    return idcs


def initialize_q_batch(X: Tensor, Y: Tensor, n: int, eta: float = 1.0) -> Tensor:
    """Pick ``n`` of the raw q-batches with probability ~ exp(eta * zscore(Y)); the best one is always kept.

    The argmax over the raw samples is the index whose parity with the reference is bit-exact
    (initializers.py:944); the multinomial draw uses torch's host RNG stream like the reference.
    """
    n_samples = X.shape[0]
    if n > n_samples:
        raise RuntimeError(f"n ({n}) cannot be larger than the number of provided samples ({n_samples})")
    if n == n_samples:
        return X
    if X.dim() != 3:
        raise UnsupportedError("initialize_q_batch on the fused path takes b x q x d raw samples (no extra batch dims)")
    return X[initialize_q_batch_indices(Y, n, eta).to(X.device)]


def gen_batch_initial_conditions(
    acq_function,
    bounds: Tensor,
    q: int,
    num_restarts: int,
    raw_samples: int,
    fixed_features: Optional[Dict[int, float]] = None,
    options: Optional[Dict[str, Union[bool, float, int]]] = None,
    inequality_constraints=None,
    equality_constraints=None,
    generator: Optional[Callable[[int, int, Optional[int]], Tensor]] = None,
    fixed_X_fantasies: Optional[Tensor] = None,
) -> Tensor:
    """``num_restarts x q x d`` starting points chosen among ``raw_samples`` Sobol q-batches by acquisition value."""
    if bool(bounds.isinf().any()):
        raise NotImplementedError(
            "Currently only finite values in `bounds` are supported for generating initial conditions for optimization."
        )
    if fixed_features or inequality_constraints or equality_constraints or fixed_X_fantasies is not None:
        raise UnsupportedError("fixed features / linear constraints / fantasies are outside the fused path")
    options = options or {}
    if options.get("sample_around_best", False) or options.get("nonnegative", False):
        raise UnsupportedError("sample_around_best / nonnegative initialisation are outside the fused path")
    seed: Optional[int] = options.get("seed")
    batch_limit: Optional[int] = options.get("init_batch_limit", options.get("batch_limit"))
    init_kwargs = {"eta": options["eta"]} if "eta" in options else {}
    device = bounds.device
    bounds_cpu = bounds.cpu()
    q = 1 if q is None else q
    effective_dim = bounds.shape[-1] * q
    factor, max_factor = 1, 5
    # base samples of the q new points: drawn on a host thread while this one draws the raw samples (both are scrambled
    # Sobol engines whose construction is independent host work)
    pending = acq_function.prefetch_base_samples(q) if hasattr(acq_function, "prefetch_base_samples") else None
    while factor < max_factor:
        with warnings.catch_warnings(record=True) as ws:
            n = raw_samples * factor
            if generator is not None:
                X_rnd = generator(n, q, seed)
            elif effective_dim <= SobolEngine.MAXDIM:
                X_rnd = draw_sobol_samples(bounds=bounds_cpu, n=n, q=q, seed=seed)
            else:
                with manual_seed(seed):
                    u = torch.rand(n, q, bounds_cpu.shape[-1], dtype=bounds.dtype)
                X_rnd = bounds_cpu[0] + (bounds_cpu[1] - bounds_cpu[0]) * u
            if pending is not None:
                pending.join()
                pending = None
            Y_rnd = evaluate_raw_samples(acq_function, X_rnd, device, batch_limit)
            ics = initialize_q_batch(X=X_rnd, Y=Y_rnd, n=num_restarts, **init_kwargs).to(device=device)
            if not any(issubclass(w.category, BadInitialCandidatesWarning) for w in ws):
                return ics
            if factor < max_factor:
                factor += 1
                if seed is not None:
                    seed += 1
    warnings.warn(
        "Unable to find non-zero acquisition function values - initial conditions are being selected randomly.",
        BadInitialCandidatesWarning,
    )
    return ics


def evaluate_raw_samples(acq_function, X_rnd: Tensor, device, batch_limit: Optional[int] = None) -> Tensor:
    """No-grad acquisition values of the raw q-batches in ``batch_limit`` chunks, returned on the host
    (initializers.py:411-423)."""
    with torch.no_grad():
        step = X_rnd.shape[0] if batch_limit is None else batch_limit
        out = [acq_function(X_rnd[i : i + step].to(device=device)).cpu() for i in range(0, X_rnd.shape[0], step)]
    _flush_warnings(acq_function)
    return torch.cat(out)


def _flush_warnings(acq_function) -> None:
    """Numerical-failure flags are read back one call late (acquisition._warn_on_failures); drain the last one."""
    check = getattr(acq_function, "check_numerics", None)
    if check is not None:
        check()


# -------------------------------------------------------------------------------------------------
# candidate generation (L-BFGS-B on the host, values/gradients from the GPU)
# -------------------------------------------------------------------------------------------------
def columnwise_clamp(X: Tensor, lower=None, upper=None, raise_on_violation: bool = False) -> Tensor:
    out = X
    if lower is not None:
        out = torch.max(out, torch.as_tensor(lower, dtype=X.dtype, device=X.device).expand_as(out))
    if upper is not None:
        out = torch.min(out, torch.as_tensor(upper, dtype=X.dtype, device=X.device).expand_as(out))
    if raise_on_violation and not bool(torch.allclose(X, out)):
        raise RuntimeError("Original value(s) are out of bounds.")
    return out


def make_scipy_bounds(X: Tensor, lower_bounds=None, upper_bounds=None) -> Optional[_sciopt.Bounds]:
    if lower_bounds is None and upper_bounds is None:
        return None

    def _expand(b, fill):
        if b is None:
            return np.full(X.numel(), fill)
        return torch.as_tensor(b, dtype=X.dtype).cpu().expand(X.shape).reshape(-1).numpy().astype(np.float64)

    return _sciopt.Bounds(lb=_expand(lower_bounds, -np.inf), ub=_expand(upper_bounds, np.inf), keep_feasible=True)


class _Timeout(Exception):
    def __init__(self, x):
        self.x = x


class HostEvaluator:
    """``f_np_wrapper`` of gen_candidates_scipy (gen.py:194-222) with the host round trip trimmed to the minimum:
    x goes through a pinned staging buffer (asynchronous H2D), the gradient and the loss come back in ONE asynchronous
    D2H into pinned memory, and the host waits once per evaluation.

    Two routes.  The general one is the reference's: ``loss = -acqf(X).sum()`` and ``torch.autograd.grad``.  For a plain
    fused acquisition function (one exact model, no pending points, float64) the objective is known in closed form, so the
    DIRECT route calls the forward and the reverse kernels itself with ``d loss / d acq = -1``, into buffers allocated once:
    no autograd graph, no per-call allocations, no reduction or concatenation launches -- the values come back next to
    the gradient and are summed on the host.  Measured at C3 (tools/e2e_profile.py): the host needs 0.29 ms to reach the
    first kernel launch on the general route and the GPU idles for it; the direct route cuts that to ~0.05 ms.  Failure
    flags come back with the same copy; an evaluation that raised any is redone on the general route, which owns the
    warnings and the reference's fallback.  ``BOTORCH_B200_DIRECT_EVAL=0`` disables the direct route."""

    def __init__(self, acquisition_function, shapeX: torch.Size, like: Tensor) -> None:
        self.acqf, self.shapeX, self.like = acquisition_function, torch.Size(shapeX), like
        numel = self.shapeX.numel()
        pin = like.is_cuda
        self.x_host = torch.empty(numel, dtype=like.dtype, pin_memory=pin)
        self.out_host = torch.empty(numel + 1, dtype=like.dtype, pin_memory=pin)
        self.calls = 0
        self.direct_calls = 0
        self._direct = None  # None: not decided yet; False: general route only; else the preallocated state

    # -- direct route ----------------------------------------------------------------------------------
    def _direct_state(self):
        if self._direct is not None:
            return self._direct
        import os
        from types import SimpleNamespace

        from . import _lib
        from .acquisition import _FusedImprovementBase

        a = self.acqf
        nei = getattr(a, "_nei", None)
        ok = (isinstance(a, _FusedImprovementBase) and nei is not None and self.like.is_cuda
              and self.like.dtype == torch.float64 and len(self.shapeX) == 3 and getattr(a, "X_pending", None) is None
              and (nei.gp.M == 1) != bool(a.model._is_ensemble) and self.like.device == nei.device
              and os.environ.get("BOTORCH_B200_DIRECT_EVAL", "1") != "0")  # fmt: skip
        if not ok:
            self._direct = False
            return False
        B, q, d = self.shapeX
        dev, numel = nei.device, self.shapeX.numel()
        out = torch.empty(numel + B, dtype=torch.float64, device=dev)  # [gradient | acquisition values]: one D2H
        M = nei.gp.M
        st = SimpleNamespace(
            X=torch.empty(B, q, d, dtype=torch.float64, device=dev), out=out, grad=out[:numel].view(1, B, q, d),
            acq=out[numel:].view(1, B), info=torch.empty(M, B, dtype=torch.int32, device=dev),
            gin=torch.full((M, B), -1.0 / M, dtype=torch.float64, device=dev),  # d(-sum_b mean_m acq) / d acq
            work=_lib.workspace(nei.workspace_bytes(B, q, True), dev),
            out_host=torch.empty(numel + B, dtype=torch.float64, pin_memory=True),
            info_host=torch.empty(M * B, dtype=torch.int32, pin_memory=True), numel=numel, M=M)  # fmt: skip
        if M > 1:  # ensemble: per-model values and gradients, mixed on the device (utils/transforms.py:289-293)
            st.mix = st.acq.view(B)
            st.acq = torch.empty(M, B, dtype=torch.float64, device=dev)
            st.grad_sum = st.grad.view(B, q, d)
            st.grad = torch.empty(M, B, q, d, dtype=torch.float64, device=dev)
        st.out_np, st.info_np, st.x_np = st.out_host.numpy(), st.info_host.numpy(), self.x_host.numpy()
        self._direct = st
        return st

    def _call_direct(self, st, x: np.ndarray):
        a = self.acqf
        nei = a._nei
        np.copyto(st.x_np, x)
        st.X.copy_(self.x_host.view(self.shapeX), non_blocking=True)
        a._set_sampler(q_in=self.shapeX[1])
        nei.forward(st.X, a._Zq, a.tau_relu, a.tau_max, True, work=st.work, out=(st.acq, st.info))
        if st.M > 1 and a._log:  # loss = -sum_b logmeanexp_m acq[m, b]:  d loss / d acq = -softmax over the models
            torch.logsumexp(st.acq, dim=0, out=st.mix)
            st.mix.sub_(math.log(st.M))
            torch.softmax(st.acq, dim=0, out=st.gin)
            st.gin.neg_()
        elif st.M > 1:
            torch.mean(st.acq, dim=0, out=st.mix)
        nei.backward(st.X, a._Zq, a.tau_relu, a.tau_max, st.gin, st.work, out=st.grad)
        if st.M > 1:
            torch.sum(st.grad, dim=0, out=st.grad_sum)
        st.out_host.copy_(st.out, non_blocking=True)
        st.info_host.copy_(st.info.view(-1), non_blocking=True)
        torch.cuda.current_stream(nei.device).synchronize()
        if st.info_np.any():  # jitter was added or a cached update failed: the general route warns / falls back
            return None
        self.direct_calls += 1
        return -float(st.out_np[st.numel :].sum()), st.out_np[: st.numel].astype(np.float64, copy=True)

    def __call__(self, x: np.ndarray, with_grad: bool = True):
        self.calls += 1
        if with_grad and not getattr(self.acqf, "_strict_numerics", False):
            st = self._direct_state()
            if st:
                res = self._call_direct(st, x)
                if res is not None:
                    return res
        self.x_host.copy_(torch.from_numpy(x))
        X = self.x_host.to(self.like.device, non_blocking=True).view(self.shapeX)
        if not with_grad:
            with torch.no_grad():
                return float((-self.acqf(X)).sum())
        X.requires_grad_(True)
        loss = (-self.acqf(X)).sum()
        (grad,) = torch.autograd.grad(loss, X)
        self.out_host.copy_(torch.cat([grad.reshape(-1), loss.detach().reshape(1)]), non_blocking=True)
        if self.like.is_cuda:
            torch.cuda.current_stream(self.like.device).synchronize()
        check = getattr(self.acqf, "check_numerics", None)
        if check is not None:
            # the failure flags of this evaluation were queued behind the kernels: they are complete now.  A failed cached
            # update is re-evaluated through the reference's fallback (standard sampling) before SciPy sees the numbers.
            check(quiet=True)
            if getattr(self.acqf, "last_call_failed", False):
                self.acqf.last_call_failed = False
                self.acqf.strict_numerics(True)
                try:
                    X = self.x_host.to(self.like.device).view(self.shapeX).requires_grad_(True)
                    loss = (-self.acqf(X)).sum()
                    (grad,) = torch.autograd.grad(loss, X)
                    self.out_host.copy_(torch.cat([grad.reshape(-1), loss.detach().reshape(1)]))
                finally:
                    self.acqf.strict_numerics(False)
        out = self.out_host.numpy()
        return float(out[-1]), out[:-1].astype(np.float64, copy=True)


def gen_candidates_scipy(
    initial_conditions: Tensor,
    acquisition_function,
    lower_bounds=None,
    upper_bounds=None,
    inequality_constraints=None,
    equality_constraints=None,
    nonlinear_inequality_constraints=None,
    options: Optional[Dict] = None,
    fixed_features: Optional[Dict[int, Optional[float]]] = None,
    timeout_sec: Optional[float] = None,
) -> Tuple[Tensor, Tensor]:
    """Maximise the summed acquisition value of all restarts jointly with L-BFGS-B (gen.py:46-298).

    Per function evaluation: host x -> device X, forward, ``torch.autograd.grad`` (the reverse CUDA kernels),
    gradient -> host.  NaN in x or the gradient raises ``RuntimeError`` as in the reference.
    """
    if inequality_constraints or equality_constraints or nonlinear_inequality_constraints or fixed_features:
        raise UnsupportedError("constraints / fixed features are outside the fused path")
    options = dict(options or {})
    options.setdefault("maxiter", 2000)
    clamped = columnwise_clamp(initial_conditions, lower_bounds, upper_bounds)
    shapeX = clamped.shape
    x0 = clamped.reshape(-1).detach().cpu().numpy().astype(np.float64)
    bounds = make_scipy_bounds(initial_conditions, lower_bounds, upper_bounds)
    with_grad = options.get("with_grad", True)

    def fun(x: np.ndarray):
        if np.isnan(x).any():
            raise RuntimeError(f"{np.isnan(x).sum()} elements of the {x.size} element array `x` are NaN.")
        if not with_grad:
            return evaluate(x, with_grad=False)
        loss, grad = evaluate(x)
        if np.isnan(grad).any():
            msg = (f"{np.isnan(grad).sum()} elements of the {x.size} element gradient array `gradf` are NaN. "
                   "This often indicates numerical issues.")  # fmt: skip
            if initial_conditions.dtype != torch.double:
                msg += " Consider using `dtype=torch.double`."
            raise RuntimeError(msg)
        return loss, grad

    evaluate = HostEvaluator(acquisition_function, shapeX, initial_conditions)
    gen_candidates_scipy.last_evaluator = evaluate  # bench.py reads the number of objective evaluations
    t0 = time.monotonic()
    user_cb = options.get("callback")

    def callback(xk):
        if timeout_sec is not None and time.monotonic() - t0 > timeout_sec:
            raise _Timeout(np.array(xk, copy=True))
        if user_cb is not None:
            user_cb(xk)

    scipy_opts = {k: v for k, v in options.items() if k not in ("method", "callback", "with_grad")}
    try:
        res = _sciopt.minimize(fun=fun, x0=x0, method=options.get("method", "L-BFGS-B"), jac=with_grad, bounds=bounds,
                               callback=callback, options=scipy_opts)  # fmt: skip
        x_final, success, status, message = res.x, res.success, res.status, res.message
    except _Timeout as e:
        x_final, success, status = e.x, False, -1
        message = f"Optimization timed out after {time.monotonic() - t0} seconds."
    if not success and "disp" not in options:  # _process_scipy_result (gen.py:458-493)
        msg = message if isinstance(message, str) else message.decode("ascii")
        if status == -1:
            warnings.warn(msg, OptimizationWarning)
        elif "ITERATIONS REACHED LIMIT" in msg or "Iteration limit reached" in msg:
            pass  # reaching maxiter is reported at INFO level only
        else:
            warnings.warn(
                f"Optimization failed within `scipy.optimize.minimize` with status {status} and message {msg}.",
                OptimizationWarning,
            )
    candidates = torch.from_numpy(x_final).to(initial_conditions).reshape(shapeX)
    clamped = columnwise_clamp(candidates, lower_bounds, upper_bounds, raise_on_violation=True)
    with torch.no_grad():
        batch_acquisition = acquisition_function(clamped)
    _flush_warnings(acquisition_function)
    return clamped, batch_acquisition


class ExpMAStoppingCriterion:
    """Stop when the exponentially weighted moving average of the objective stalls (optim/stopping.py:30-127)."""

    def __init__(self, maxiter: int = 10000, minimize: bool = True, n_window: int = 10, eta: float = 1.0,
                 rel_tol: float = 1e-5) -> None:  # fmt: skip
        self.maxiter, self.minimize, self.n_window, self.rel_tol = maxiter, minimize, n_window, rel_tol
        self.iter = 0
        weights = torch.exp(torch.linspace(-eta, 0, self.n_window))
        self.weights = weights / weights.sum()
        self._prev_fvals = None

    def evaluate(self, fvals: Tensor) -> bool:
        self.iter += 1
        if self.iter == self.maxiter:
            return True
        if self._prev_fvals is None:
            self._prev_fvals = fvals.unsqueeze(0)
        else:
            self._prev_fvals = torch.cat([self._prev_fvals[-self.n_window :], fvals.unsqueeze(0)])
        if self._prev_fvals.size(0) < self.n_window + 1:
            return False
        weights = self.weights.to(fvals)
        if self._prev_fvals.ndim > 1:
            weights = weights.unsqueeze(-1)
        prev_ma = (self._prev_fvals[:-1] * weights).sum(dim=0)
        ma = (self._prev_fvals[1:] * weights).sum(dim=0)
        rel_delta = (prev_ma - ma) / prev_ma.abs()
        if not self.minimize:
            rel_delta = -rel_delta
        return bool(torch.max(rel_delta) < self.rel_tol)


def gen_candidates_torch(
    initial_conditions: Tensor,
    acquisition_function,
    lower_bounds=None,
    upper_bounds=None,
    optimizer=torch.optim.Adam,
    options: Optional[Dict] = None,
    callback: Optional[Callable] = None,
    fixed_features: Optional[Dict[int, Optional[float]]] = None,
    timeout_sec: Optional[float] = None,
) -> Tuple[Tensor, Tensor]:
    """Maximise the summed acquisition value with a ``torch.optim`` optimiser (gen.py:301-424).

    The iterate never leaves the device: per iteration one fused forward, one fused backward and the optimiser's
    elementwise update -- no SciPy round trip.  Stopping: ``ExpMAStoppingCriterion``.
    """
    if fixed_features:
        raise UnsupportedError("fixed features are outside the fused path")
    start_time = time.monotonic()
    options = dict(options or {})
    clamped_candidates = columnwise_clamp(initial_conditions, lower_bounds, upper_bounds).requires_grad_(True)
    _optimizer = optimizer(params=[clamped_candidates], lr=options.get("lr", 0.025))
    stop_keys = ("maxiter", "minimize", "n_window", "eta", "rel_tol")
    stopping_criterion = ExpMAStoppingCriterion(**{k: v for k, v in options.items() if k in stop_keys})
    i, stop = 0, False
    while not stop:
        i += 1
        with torch.no_grad():
            X = columnwise_clamp(clamped_candidates, lower_bounds, upper_bounds).requires_grad_(True)
        loss = -acquisition_function(X).sum()
        grad = torch.autograd.grad(loss, X)[0]
        if callback:
            callback(i, loss, grad)

        def assign_grad():
            _optimizer.zero_grad()
            clamped_candidates.grad = grad
            return loss

        _optimizer.step(assign_grad)
        stop = stopping_criterion.evaluate(fvals=loss.detach())
        if timeout_sec is not None and time.monotonic() - start_time > timeout_sec:
            stop = True
    clamped_candidates = columnwise_clamp(clamped_candidates.detach(), lower_bounds, upper_bounds)
    with torch.no_grad():
        batch_acquisition = acquisition_function(clamped_candidates)
    _flush_warnings(acquisition_function)
    return clamped_candidates, batch_acquisition


def gen_candidates_lbfgs(
    initial_conditions: Tensor,
    acquisition_function,
    lower_bounds=None,
    upper_bounds=None,
    options: Optional[Dict] = None,
    fixed_features: Optional[Dict[int, Optional[float]]] = None,
    timeout_sec: Optional[float] = None,
) -> Tuple[Tensor, Tensor]:
    """Multi-start maximisation with a device-resident batched projected L-BFGS (``csrc/lbfgs.cu``).

    Same contract as ``gen_candidates_scipy`` / ``gen_candidates_torch`` (gen.py:46-57): B x q x d starting points in,
    (candidates, acquisition values) out.  The reference runs ONE SciPy L-BFGS-B over all B*q*d variables; the
    objective is a sum over restarts (gen.py:209), so here every restart is its own box-constrained problem with its
    own curvature model, and the iterate never leaves the GPU: per round one ``bo_lbfgs_propose``, one fused
    forward + backward of the acquisition function, one ``bo_lbfgs_accept`` -- no host synchronisation except a
    convergence poll every ``check_every`` rounds.  Options: maxiter (2000), maxcor (10), ftol (2.22e-9), gtol (1e-5)
    as in SciPy's L-BFGS-B, c1 (1e-4), check_every (8), compact (True: shrink the batch as restarts stop).
    """
    from . import _lib  # deferred: importing the package must not need the library

    if fixed_features:
        raise UnsupportedError("fixed features are outside the fused path")
    options = dict(options or {})
    start_time = time.monotonic()
    X0 = columnwise_clamp(initial_conditions, lower_bounds, upper_bounds).detach()
    dev = X0.device
    if dev.type != "cuda":
        raise _lib.BoError("gen_candidates_lbfgs keeps its state on a CUDA device; there is no CPU path")
    shapeX = X0.shape
    B, D = shapeX[0], int(torch.Size(shapeX[1:]).numel())
    m = int(options.get("maxcor", 10))
    maxiter = int(options.get("maxiter", 2000))
    ftol, gtol = float(options.get("ftol", 2.220446049250313e-09)), float(options.get("gtol", 1e-5))
    c1, check_every = float(options.get("c1", 1e-4)), int(options.get("check_every", 8))
    compact = bool(options.get("compact", True))

    def bound(b, fill):
        t = torch.full(shapeX[1:], fill, dtype=torch.float64, device=dev)
        if b is not None:
            t = torch.as_tensor(b, dtype=torch.float64, device=dev).expand(shapeX[1:]).clone()
        return t.reshape(-1).contiguous()

    f64 = lambda *s_: torch.zeros(*s_, dtype=torch.float64, device=dev)
    i32 = lambda *s_: torch.zeros(*s_, dtype=torch.int32, device=dev)
    bufs = dict(x=X0.to(torch.float64).reshape(B, D).contiguous().clone(), g=f64(B, D), f=f64(B), x_prev=f64(B, D),
                g_prev=f64(B, D), S=f64(B, m, D), Y=f64(B, m, D), rho=f64(B, m), hist_count=i32(B), hist_head=i32(B),
                p=f64(B, D), alpha=f64(B), x_trial=f64(B, D), state=i32(B), lower=bound(lower_bounds, -float("inf")),
                upper=bound(upper_bounds, float("inf")), n_iter=i32(B))  # fmt: skip
    lib, ctx = _lib.lib(), _lib.ctx(dev)
    shared = ("lower", "upper")  # not per restart

    def bind(b: dict):
        return _lib.BoLBFGS(B=b["x"].shape[0], D=D, m=m, **{k: _lib.ptr(v) for k, v in b.items()})

    def evaluate(x_flat: Tensor):
        Bc = x_flat.shape[0]
        X = x_flat.view(Bc, *shapeX[1:]).to(initial_conditions.dtype).requires_grad_(True)
        acq = acquisition_function(X)
        (grad,) = torch.autograd.grad(acq.sum(), X)
        return (-acq.detach()).to(torch.float64).contiguous(), (-grad).to(torch.float64).reshape(Bc, D).contiguous()

    st = bind(bufs)
    f0, g0 = evaluate(bufs["x"])
    _lib.check(lib.bo_lbfgs_init(ctx, C.byref(st), _lib.ptr(f0), _lib.ptr(g0), _lib.stream_ptr(dev)))
    # Restarts stop at very different times.  Every poll the batch is compacted to the next size of a fixed ladder
    # (B, B/2, B/4, ...: few distinct shapes, so the captured graphs are reused) that still holds all running restarts;
    # stopped ones leave with their final point.
    ladder = sorted({max(int(math.ceil(B / 2**k)), min(B, 16)) for k in range(0, 12)}, reverse=True)
    final_x = bufs["x"].clone()
    final_iter = torch.zeros(B, dtype=torch.int32, device=dev)
    rows = torch.arange(B, device=dev)  # global restart index of every row of the current batch
    rounds, evals = 0, B
    max_rounds = maxiter * 4  # accepted steps are bounded per restart by maxiter; rounds include rejected trials
    while rounds < max_rounds:
        _lib.check(lib.bo_lbfgs_propose(ctx, C.byref(st), _lib.stream_ptr(dev)))
        ft, gt = evaluate(bufs["x_trial"])
        _lib.check(lib.bo_lbfgs_accept(ctx, C.byref(st), _lib.ptr(ft), _lib.ptr(gt), c1, ftol, gtol, maxiter,
                                       _lib.stream_ptr(dev)))  # fmt: skip
        rounds += 1
        evals += rows.numel()
        if rounds % check_every == 0:
            running = bufs["state"] != 2
            n_run = int(running.sum())  # the one host synchronisation per poll
            if n_run == 0:
                break
            if timeout_sec is not None and time.monotonic() - start_time > timeout_sec:
                break
            target = min(s_ for s_ in ladder if s_ >= n_run)
            if compact and target < rows.numel():
                final_x[rows] = bufs["x"]
                final_iter[rows] = bufs["n_iter"]
                keep = torch.cat([running.nonzero().view(-1), (~running).nonzero().view(-1)])[:target]
                bufs = {k: (v if k in shared else v[keep].contiguous()) for k, v in bufs.items()}
                rows = rows[keep]
                st = bind(bufs)
    final_x[rows] = bufs["x"]
    final_iter[rows] = bufs["n_iter"]
    bufs = {"x": final_x, "n_iter": final_iter}
    candidates = bufs["x"].view(shapeX).to(initial_conditions.dtype)
    candidates = columnwise_clamp(candidates, lower_bounds, upper_bounds, raise_on_violation=True)
    with torch.no_grad():
        batch_acquisition = acquisition_function(candidates)
    _flush_warnings(acquisition_function)
    gen_candidates_lbfgs.last_stats = {"rounds": rounds, "accepted_steps": bufs["n_iter"].clone(),
                                       "restart_evaluations": evals}
    return candidates, batch_acquisition


def get_best_candidates(batch_candidates: Tensor, batch_values: Tensor) -> Tensor:
    best = torch.argmax(batch_values.view(-1), dim=0)
    return batch_candidates[best]


# -------------------------------------------------------------------------------------------------
# optimize_acqf
# -------------------------------------------------------------------------------------------------
def optimize_acqf(
    acq_function,
    bounds: Tensor,
    q: int,
    num_restarts: int,
    raw_samples: Optional[int] = None,
    options: Optional[Dict[str, Union[bool, float, int, str]]] = None,
    inequality_constraints=None,
    equality_constraints=None,
    nonlinear_inequality_constraints=None,
    fixed_features: Optional[Dict[int, float]] = None,
    post_processing_func: Optional[Callable[[Tensor], Tensor]] = None,
    batch_initial_conditions: Optional[Tensor] = None,
    return_best_only: bool = True,
    gen_candidates: Optional[Callable] = None,
    sequential: bool = False,
    *,
    ic_generator: Optional[Callable] = None,
    timeout_sec: Optional[float] = None,
    return_full_tree: bool = False,
    retry_on_optimization_warning: bool = True,
    **ic_gen_kwargs,
) -> Tuple[Tensor, Tensor]:
    """Multi-start maximisation of ``acq_function`` over the box ``bounds`` (optimize.py:397-543).

Joint q-batch optimisation: raw samples -> initial conditions -> ``batch_limit`` chunks of restarts, each optimised
    jointly by ``gen_candidates`` (default L-BFGS-B), one retry on an ``OptimizationWarning``, best restart returned.
    ``sequential=True`` picks the q points greedily one at a time (optimize.py:202-243).
    """
    if inequality_constraints or equality_constraints or nonlinear_inequality_constraints or fixed_features:
        raise UnsupportedError("constraints / fixed features are outside the fused path (SURVEY.md section 8)")
    if sequential and q > 1:
        # _optimize_acqf_sequential_q (optimize.py:202-243): q greedy single-point optimisations, each conditioned on the
        # points chosen so far through X_pending (they enter the fused kernels as extra columns of the q-batch)
        if not return_best_only:
            raise NotImplementedError("`return_best_only=False` only supported for joint optimization.")
        if batch_initial_conditions is not None:
            raise UnsupportedError("`batch_initial_conditions` is not supported for sequential optimization.")
        candidate_list, acq_value_list = [], []
        base_X_pending = acq_function.X_pending
        step_timeout = timeout_sec / q if timeout_sec is not None else None
        candidates = None
        for _ in range(q):
            candidate, acq_value = optimize_acqf(
                acq_function, bounds, q=1, num_restarts=num_restarts, raw_samples=raw_samples, options=options,
                post_processing_func=post_processing_func, return_best_only=True, gen_candidates=gen_candidates,
                sequential=False, ic_generator=ic_generator, timeout_sec=step_timeout,
                retry_on_optimization_warning=retry_on_optimization_warning, **ic_gen_kwargs)  # fmt: skip
            candidate_list.append(candidate)
            acq_value_list.append(acq_value)
            candidates = torch.cat(candidate_list, dim=-2)
            acq_function.set_X_pending(
                torch.cat([base_X_pending, candidates], dim=-2) if base_X_pending is not None else candidates
            )
        acq_function.set_X_pending(base_X_pending)
        return candidates, torch.stack(acq_value_list)
    if not (bounds.ndim == 2 and bounds.shape[0] == 2):
        raise ValueError(f"bounds should be a `2 x d` tensor, current shape: {list(bounds.shape)}.")
    if batch_initial_conditions is None and raw_samples is None and ic_generator is None:
        raise ValueError("Must specify `raw_samples` when `batch_initial_conditions` is None`.")
    gen_candidates = gen_candidates or gen_candidates_scipy
    ic_generator = ic_generator or gen_batch_initial_conditions
    options = options or {}

    def initial_conditions() -> Tensor:
        return ic_generator(acq_function=acq_function, bounds=bounds, q=q, num_restarts=num_restarts,
                            raw_samples=raw_samples, fixed_features=None,
                            options={k: v for k, v in options.items() if k in INIT_OPTION_KEYS}, **ic_gen_kwargs)  # fmt: skip

    provided_ics = batch_initial_conditions is not None
    ics = batch_initial_conditions if provided_ics else initial_conditions()
    batch_limit: int = options.get("batch_limit", num_restarts)
    gen_options = {k: v for k, v in options.items() if k not in INIT_OPTION_KEYS}
    start = time.monotonic()

    def run(ics: Tensor):
        cands, vals, ws_all = [], [], []
        chunks = ics.split(batch_limit)
        for chunk in chunks:
            remaining = None if timeout_sec is None else max(timeout_sec - (time.monotonic() - start), 1e-3) / len(chunks)
            with warnings.catch_warnings(record=True) as ws:
                warnings.simplefilter("always", category=OptimizationWarning)
                c, v = gen_candidates(chunk, acq_function, lower_bounds=bounds[0], upper_bounds=bounds[1],
                                      options=gen_options, timeout_sec=remaining)  # fmt: skip
            ws_all += [w for w in ws if issubclass(w.category, OptimizationWarning)]
            for w in ws:
                if not issubclass(w.category, OptimizationWarning):
                    warnings.warn_explicit(w.message, w.category, w.filename, w.lineno)
            cands.append(c)
            vals.append(v)
        return torch.cat(cands), torch.cat(vals), ws_all

    batch_candidates, batch_acq_values, ws = run(ics)
    if ws and retry_on_optimization_warning and not provided_ics:
        warnings.warn(
            "Optimization failed in `gen_candidates_scipy` with the following warning(s):\n"
            f"{[w.message for w in ws]}\nTrying again with a new set of initial conditions.",
            RuntimeWarning,
        )
        batch_candidates, batch_acq_values, ws = run(initial_conditions())
        if ws:
            warnings.warn("Optimization failed on the second try, after generating a new set of initial conditions.",
                          RuntimeWarning)  # fmt: skip
    elif ws:
        for w in ws:
            warnings.warn_explicit(w.message, w.category, w.filename, w.lineno)
    if post_processing_func is not None:
        batch_candidates = post_processing_func(batch_candidates)
        with torch.no_grad():
            batch_acq_values = acq_function(batch_candidates)
    if return_best_only:
        best = torch.argmax(batch_acq_values.view(-1), dim=0)
        batch_candidates = batch_candidates[best]
        batch_acq_values = batch_acq_values[best]
    return batch_candidates, batch_acq_values
