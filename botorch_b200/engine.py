"""Device state of a fitted GP and of a qLogNEI baseline, driving libbotorch_b200 through its C ABI.

Host-side logic only: buffer ownership, the psd_safe_cholesky retry schedule
(linear_operator 0.5.2 semantics; botorch/__init__.py:47 sets 6 tries), warnings and errors.
Every number is produced by the CUDA library.
"""
from __future__ import annotations

import ctypes as C
import os
import warnings
from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _lib
from ._lib import BoGP, BoNEI, BoPaths, check, kp, pad8, ptr
from .exceptions import NanError, NotPSDError, NumericalWarning

REDUCE_LOG, REDUCE_PLAIN, REDUCE_LOG_SOFT = 0, 1, 2  # BO_REDUCE_* of include/botorch_b200.h (LOG_SOFT: fat = False)
MAX_Q = 32  # BO_MAX_Q: largest q-batch of the per-candidate kernels

KERNEL_IDS = {"rbf": 0, "matern52": 1}
CHOLESKY_MAX_TRIES = 6  # botorch/__init__.py:47
CHOLESKY_JITTER = 1e-8  # linear_operator settings.cholesky_jitter for float64


def _f64(t, device) -> Tensor:
    return torch.as_tensor(t, dtype=torch.float64, device=device).contiguous()


def _retry_with_jitter(run, M: int, device, what: str) -> Tensor:
    """psd_safe_cholesky: plain attempt, then jitter 1e-8*10**i on the members that failed."""
    jitter = torch.zeros(M, dtype=torch.float64, device=device)
    info = run(None)
    if not bool(info.any()):
        return jitter
    prev_fail = info > 0
    for i in range(CHOLESKY_MAX_TRIES):
        new = CHOLESKY_JITTER * (10**i)
        jitter = torch.where(prev_fail, torch.full_like(jitter, new), jitter)
        warnings.warn(f"{what} not p.d., added jitter of {new:.1e} to the diagonal", NumericalWarning)
        info = run(jitter)
        prev_fail = info > 0
        if not bool(prev_fail.any()):
            return jitter
    raise NotPSDError(f"{what} not positive definite after repeatedly adding jitter up to {new:.1e}.")


class GPState:
    """Training-set factors of an exact GP (one or M batched hyperparameter sets) resident in HBM."""

    def __init__(
        self,
        train_X: Tensor,
        targets: Tensor,
        kernel: str,
        lengthscale: Tensor,
        noise: Tensor,
        mean: Tensor,
        outputscale: Optional[Tensor] = None,
        y_mean: float = 0.0,
        y_std: float = 1.0,
        setup: bool = True,
        noise_vec: Optional[Tensor] = None,
    ) -> None:
        dev = train_X.device
        self.device = dev
        self.ctx = _lib.ctx(dev)
        self.kernel = kernel
        self.train_X = _f64(train_X, dev)
        self.n, self.d = self.train_X.shape
        ls = _f64(lengthscale, dev).reshape(-1, self.d)
        self.M = ls.shape[0]
        self.batched = _f64(lengthscale, dev).dim() > 1
        self.inv_ls = (1.0 / ls).contiguous()
        self.noise = _f64(noise, dev).reshape(-1).expand(self.M).contiguous()
        self.mean = _f64(mean, dev).reshape(-1).expand(self.M).contiguous()
        self.outputscale = None if outputscale is None else _f64(outputscale, dev).reshape(-1).expand(self.M).contiguous()
        self.targets = _f64(targets, dev).reshape(-1)
        # fixed per-observation noise [n] (shared by the M models): K + diag(noise_vec); `noise` stays the level that
        # observation_noise=True adds at test points
        self.noise_vec = None if noise_vec is None else _f64(noise_vec, dev).reshape(1, -1).expand(self.M, -1).contiguous()
        if self.noise_vec is not None and self.noise_vec.shape[1] != self.train_X.shape[0]:
            raise ValueError(f"noise_vec needs one entry per training point, got {tuple(self.noise_vec.shape)}")
        self.center = self.train_X.mean(dim=0).contiguous()
        self.y_mean, self.y_std = float(y_mean), float(y_std)
        for name, t in (("train_X", self.train_X), ("targets", self.targets), ("lengthscale", ls), ("noise", self.noise)):
            if bool(torch.isnan(t).any()):
                raise NanError(f"{name} contains NaN")
        n, M, d = self.n, self.M, self.d
        self.np_, self.n1p, self.Kp = pad8(n), pad8(n + 1), kp(d)
        z = lambda *s: torch.zeros(*s, dtype=torch.float64, device=dev)
        self.aug_train = z(M, n, self.Kp)
        self.linv_aug = z(M, n + 1, self.np_)
        self.linv_aug_t = z(M, n, self.n1p)
        self.chol = z(M, n, self.np_)
        self.xt_aug = z(d + 1, self.np_)
        self.info = torch.zeros(M, dtype=torch.int32, device=dev)
        self.c = BoGP(
            kernel_id=KERNEL_IDS[kernel], n=n, d=d, M=M, y_mean=self.y_mean, y_std=self.y_std,
            train_X=ptr(self.train_X), center=ptr(self.center), inv_ls=ptr(self.inv_ls),
            outputscale=ptr(self.outputscale), noise=ptr(self.noise), mean=ptr(self.mean),
            targets=ptr(self.targets), jitter=None,
            aug_train=ptr(self.aug_train), linv_aug=ptr(self.linv_aug), linv_aug_t=ptr(self.linv_aug_t),
            chol=ptr(self.chol), xt_aug=ptr(self.xt_aug), noise_vec=ptr(self.noise_vec),
        )  # fmt: skip
        self.jitter = torch.zeros(M, dtype=torch.float64, device=dev)
        if not setup:  # factors arrive by broadcast from the rank that ran the setup (distributed.py)
            return
        lib = _lib.lib()
        work = _lib.workspace(lib.bo_gp_setup_workspace_bytes(C.byref(self.c)), dev)

        def run(jitter: Optional[Tensor]) -> Tensor:
            self._jitter = jitter
            self.c.jitter = ptr(jitter)
            check(lib.bo_gp_setup(self.ctx, C.byref(self.c), work.data_ptr(), work.numel(), ptr(self.info),
                                  _lib.stream_ptr(dev)))  # fmt: skip
            return self.info.clone()

        self.jitter = _retry_with_jitter(run, M, dev, "K + noise*I")

    def factor_buffers(self):
        """Device buffers bo_gp_setup fills: what the setup rank broadcasts to its peers."""
        return [self.aug_train, self.linv_aug, self.linv_aug_t, self.chol, self.xt_aug, self.jitter]

    # -- factors exposed for tests / callers ------------------------------------------------------
    @property
    def L(self) -> Tensor:
        return self.chol[:, :, : self.n]

    @property
    def L_inv(self) -> Tensor:
        return self.linv_aug[:, : self.n, : self.n]

    @property
    def alpha(self) -> Tensor:
        return self.linv_aug[:, self.n, : self.n]

    # -- posterior --------------------------------------------------------------------------------
    def posterior_joint(self, X: Tensor) -> Tuple[Tensor, Tensor]:
        """X: B x q x d -> mean [M x B x q], cov [M x B x q x q] on the outcome scale."""
        X = _f64(X, self.device)
        B, q, _ = X.shape
        lib = _lib.lib()
        if q > MAX_Q:  # large joint sets (e.g. all of X_baseline): GEMM-shaped, one t-batch at a time, no gradient
            ldq = pad8(q)
            mean = torch.empty(B, self.M, q, dtype=torch.float64, device=self.device)
            cov = torch.empty(B, self.M, q, ldq, dtype=torch.float64, device=self.device)
            work = _lib.workspace(lib.bo_posterior_joint_large_workspace_bytes(C.byref(self.c), q), self.device)
            for b in range(B):
                check(lib.bo_posterior_joint_large(self.ctx, C.byref(self.c), q, ptr(X[b]), ptr(mean[b]), ptr(cov[b]),
                                                   work.data_ptr(), work.numel(), _lib.stream_ptr(self.device)))  # fmt: skip
            return mean.transpose(0, 1), cov[..., :q].transpose(0, 1)
        mean = torch.empty(self.M, B, q, dtype=torch.float64, device=self.device)
        cov = torch.empty(self.M, B, q, q, dtype=torch.float64, device=self.device)
        work = _lib.workspace(lib.bo_posterior_workspace_bytes(C.byref(self.c), B * q), self.device)
        check(lib.bo_posterior_joint(self.ctx, C.byref(self.c), B, q, ptr(X), ptr(mean), ptr(cov), work.data_ptr(),
                                     work.numel(), _lib.stream_ptr(self.device)))  # fmt: skip
        return mean, cov

    def posterior_mean_var(self, Xc: Tensor, work: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
        """Xc: Nc x d -> marginal mean, var [M x Nc]."""
        Xc = _f64(Xc, self.device)
        Nc = Xc.shape[0]
        lib = _lib.lib()
        mean = torch.empty(self.M, Nc, dtype=torch.float64, device=self.device)
        var = torch.empty(self.M, Nc, dtype=torch.float64, device=self.device)
        if work is None:
            work = _lib.workspace(lib.bo_posterior_workspace_bytes(C.byref(self.c), Nc), self.device)
        check(lib.bo_posterior_mean_var(self.ctx, C.byref(self.c), Nc, ptr(Xc), ptr(mean), ptr(var), work.data_ptr(),
                                        work.numel(), _lib.stream_ptr(self.device)))  # fmt: skip
        return mean, var

    def posterior_workspace(self, Nc: int) -> Tensor:
        return _lib.workspace(_lib.lib().bo_posterior_workspace_bytes(C.byref(self.c), Nc), self.device)

    def kernel_matrix(self, Xc: Tensor) -> Tensor:
        """K(train_X, Xc) per model: [M x n x Nc] (parity-test hook)."""
        Xc = _f64(Xc, self.device)
        Nc = Xc.shape[0]
        out = torch.empty(self.M, self.n, pad8(Nc), dtype=torch.float64, device=self.device)
        work = _lib.workspace(8 * self.M * self.Kp * pad8(Nc) + 512, self.device)
        check(_lib.lib().bo_kernel_matrix(self.ctx, C.byref(self.c), Nc, ptr(Xc), ptr(out), work.data_ptr(),
                                          work.numel(), _lib.stream_ptr(self.device)))  # fmt: skip
        return out[:, :, :Nc]


class MLLState:
    """Exact marginal log likelihood + hyperparameter gradient on the device (``bo_mll_value_grad``).

    Owns one ``GPState`` whose hyperparameter buffers are overwritten in place per evaluation (no reallocation inside the
    L-BFGS-B loop of ``fit_gpytorch_mll``); each evaluation leaves that state factored for its hyperparameters.
    """

    def __init__(self, train_X: Tensor, targets: Tensor, kernel: str, with_outputscale: bool, M: int = 1,
                 noise_vec: Optional[Tensor] = None) -> None:  # fmt: skip
        dev = train_X.device
        d = train_X.shape[-1]
        one = torch.ones(M, dtype=torch.float64, device=dev)
        self.gp = GPState(train_X, targets, kernel, torch.ones(M, d, dtype=torch.float64, device=dev), one, 0.0 * one,
                          one.clone() if with_outputscale else None, setup=False, noise_vec=noise_vec)  # fmt: skip
        lib = _lib.lib()
        self.work = _lib.workspace(lib.bo_mll_workspace_bytes(C.byref(self.gp.c)), dev)
        self.value = torch.zeros(M, dtype=torch.float64, device=dev)
        self.grad = torch.zeros(M, d + 3, dtype=torch.float64, device=dev)
        self.evaluations = 0

    def evaluate(self, lengthscale: Tensor, noise: Optional[Tensor], mean: Tensor, outputscale: Optional[Tensor] = None,
                 retry: bool = True):
        """-> (value [M], grad [M x (d+3)]: lengthscale (d), outputscale, noise, mean).  Raises NotPSDError after the
        psd_safe_cholesky jitter schedule is exhausted; with ``retry=False`` there is one plain attempt and the per-member
        outcome is left in ``self.info`` (0 ok, >0 first bad pivot, <0 NaN/Inf) for the caller to act on."""
        gp = self.gp
        gp.inv_ls.copy_(1.0 / _f64(lengthscale, gp.device).reshape(gp.M, gp.d))
        if noise is not None:
            gp.noise.copy_(_f64(noise, gp.device).reshape(-1).expand(gp.M))
        gp.mean.copy_(_f64(mean, gp.device).reshape(-1).expand(gp.M))
        if gp.outputscale is not None and outputscale is not None:
            gp.outputscale.copy_(_f64(outputscale, gp.device).reshape(-1).expand(gp.M))
        lib = _lib.lib()

        def run(jitter: Optional[Tensor]) -> Tensor:
            gp.c.jitter = ptr(jitter)
            self._jitter = jitter
            check(lib.bo_mll_value_grad(gp.ctx, C.byref(gp.c), self.work.data_ptr(), self.work.numel(), ptr(self.value),
                                        ptr(self.grad), ptr(gp.info), _lib.stream_ptr(gp.device)))  # fmt: skip
            return gp.info.clone()

        self.evaluations += 1
        if not retry:
            self.info = run(None)
            return self.value, self.grad
        gp.jitter = _retry_with_jitter(run, gp.M, gp.device, "K + noise*I")
        return self.value, self.grad


class PathState:
    """S pathwise (Matheron) posterior samples of a single-model GP: random-feature prior paths + the exact data update
    (``bo_paths_setup``), evaluated at any number of points with ``bo_paths_eval``.  The random draws are inputs."""

    def __init__(self, gp: GPState, W: Tensor, w: Tensor, eps: Tensor) -> None:
        if gp.M != 1:
            raise ValueError("pathwise samples are defined for single-model GPs (M = 1)")
        dev = gp.device
        self.gp = gp
        self.W, self.w, self.eps = _f64(W, dev), _f64(w, dev), _f64(eps, dev)
        half, d = self.W.shape
        self.S, self.F = self.w.shape
        if d != gp.d or self.F != 2 * half or self.eps.shape != (self.S, gp.n):
            raise ValueError(f"inconsistent draws: W {tuple(self.W.shape)}, w {tuple(self.w.shape)}, eps {tuple(self.eps.shape)}")
        z = lambda *s: torch.zeros(*s, dtype=torch.float64, device=dev)
        self.w_aug, self.w_pad, self.v = z(half, gp.Kp), z(self.S, pad8(self.F)), z(self.S, gp.np_)
        self.w_t, self.v_t = z(self.F, pad8(self.S)), z(gp.n, pad8(self.S))
        self.c = BoPaths(S=self.S, F=self.F, W=ptr(self.W), w=ptr(self.w), eps=ptr(self.eps), w_aug=ptr(self.w_aug),
                         w_pad=ptr(self.w_pad), v=ptr(self.v), w_t=ptr(self.w_t), v_t=ptr(self.v_t))  # fmt: skip
        lib = _lib.lib()
        work = _lib.workspace(lib.bo_paths_setup_workspace_bytes(C.byref(gp.c), C.byref(self.c)), dev)
        check(lib.bo_paths_setup(gp.ctx, C.byref(gp.c), C.byref(self.c), work.data_ptr(), work.numel(), _lib.stream_ptr(dev)))
        self._work: Optional[Tensor] = None
        self._work_n = 0

    def evaluate(self, Xc: Tensor) -> Tensor:
        """Xc: Nc x d -> path values [S x Nc] on the outcome scale."""
        gp = self.gp
        Xc = _f64(Xc, gp.device)
        Nc = Xc.shape[0]
        lib = _lib.lib()
        if self._work is None or self._work_n < Nc:
            self._work = _lib.workspace(lib.bo_paths_eval_workspace_bytes(C.byref(gp.c), C.byref(self.c), Nc), gp.device)
            self._work_n = Nc
        out = torch.empty(self.S, pad8(Nc), dtype=torch.float64, device=gp.device)
        check(lib.bo_paths_eval(gp.ctx, C.byref(gp.c), C.byref(self.c), Nc, ptr(Xc), ptr(out), self._work.data_ptr(),
                                self._work.numel(), _lib.stream_ptr(gp.device)))  # fmt: skip
        return out[:, :Nc]

    def gradient(self, Xc: Tensor, out_bar: Tensor) -> Tensor:
        """d(sum out_bar * paths(Xc)) / dXc: Xc [Nc x d], out_bar [S x Nc] -> [Nc x d] (``bo_paths_grad``)."""
        gp = self.gp
        Xc = _f64(Xc, gp.device)
        Nc = Xc.shape[0]
        lib = _lib.lib()
        bar = torch.zeros(self.S, pad8(Nc), dtype=torch.float64, device=gp.device)
        bar[:, :Nc] = out_bar
        work = _lib.workspace(lib.bo_paths_grad_workspace_bytes(C.byref(gp.c), C.byref(self.c), Nc), gp.device)
        grad = torch.empty(Nc, gp.d, dtype=torch.float64, device=gp.device)
        check(lib.bo_paths_grad(gp.ctx, C.byref(gp.c), C.byref(self.c), Nc, ptr(Xc), ptr(bar), ptr(grad), work.data_ptr(),
                                work.numel(), _lib.stream_ptr(gp.device)))  # fmt: skip
        return grad


class _PathEval(torch.autograd.Function):
    """paths(X) with the hand-written reverse pass (features and covariance are recomputed, nothing but X is saved)."""

    @staticmethod
    def forward(ctx, X: Tensor, ps: "PathState"):
        Xd = _f64(X.detach(), ps.gp.device)
        ctx.ps, ctx.in_dtype = ps, X.dtype
        ctx.save_for_backward(Xd)
        return ps.evaluate(Xd)

    @staticmethod
    def backward(ctx, out_bar: Tensor):
        (Xd,) = ctx.saved_tensors
        return ctx.ps.gradient(Xd, out_bar.to(torch.float64)).to(ctx.in_dtype), None


def path_values(ps: "PathState", X: Tensor) -> Tensor:
    """X: N x d -> [S x N]; differentiable w.r.t. X."""
    if torch.is_grad_enabled() and X.requires_grad:
        return _PathEval.apply(X, ps)
    return ps.evaluate(X.detach())


# CUDA graphs pay only where a step is launch-bound (C1 / C2: 0.1 ms for ~11 launches) AND the shape recurs often enough
# to amortise the capture (tools/graph_cost.py: 30-46 ms for the first capture of a process, 2-4 ms afterwards, <= 20 us
# saved per replay).  At C3 a step is 1.5 ms of kernels that the host enqueues well ahead of the GPU, so shapes whose
# workspace exceeds GRAPH_MAX_WORKSPACE_BYTES stay eager; smaller ones are captured at their GRAPH_MIN_USES-th use
# (BOTORCH_B200_GRAPH_AFTER overrides, BOTORCH_B200_GRAPHS=0 disables).
GRAPH_MAX_WORKSPACE_BYTES = 16 << 20
GRAPH_MIN_USES = 64


def _local_derivs_default(reduction: int) -> int:
    """bo_nei.local_derivs: with the log reductions a gradient-carrying forward leaves d qmax_s / d f in the sample rows
    (the reverse kernel then only scales them).  ``keep_samples()`` switches it off for callers that read the samples.
    BOTORCH_B200_LOCAL_DERIVS=0 disables it for A/B measurements."""
    return int(int(reduction) != REDUCE_PLAIN and os.environ.get("BOTORCH_B200_LOCAL_DERIVS", "1") != "0")


class NEIState:
    """Baseline factors of qLogNEI (V_b, L_b and its inverse, base samples, best_f) and the fused fwd/bwd."""

    def __init__(self, gp: GPState, X_baseline: Tensor, Z_b, setup: bool = True,
                 reduction: int = REDUCE_LOG, obj_weight: float = 1.0, forced_jitter: Optional[float] = None,
                 S: Optional[int] = None) -> None:  # fmt: skip
        """``Z_b``: the S x n_b base samples of the baseline points, or a callable returning them (then ``S`` is required):
        the factors that do not read them are enqueued first (``bo_nei.setup_phase`` 1) and the callable runs on the host
        while the device works -- drawing them is a scrambled Sobol engine of dimension n_b, ~10 ms at n_b = 2000, about
        what the device needs for the factors."""
        dev = gp.device
        self.gp = gp
        self.device = dev
        self.X_b = _f64(X_baseline, dev)
        z_producer = Z_b if callable(Z_b) else None
        if z_producer is not None:
            if S is None:
                raise ValueError("S is required when the base samples are produced lazily")
            self.Z_b = torch.empty(int(S), self.X_b.shape[0], dtype=torch.float64, device=dev)
        else:
            self.Z_b = _f64(Z_b, dev)
        self.S, self.n_b = self.Z_b.shape
        if self.X_b.shape != (self.n_b, gp.d):
            raise ValueError(f"X_baseline {tuple(self.X_b.shape)} does not match base samples {tuple(self.Z_b.shape)}")
        self.is_train = self.X_b.shape == gp.train_X.shape and bool(torch.equal(self.X_b, gp.train_X))
        n, M, nb, S = gp.n, gp.M, self.n_b, self.S
        nbp, Sp = pad8(nb), pad8(S)
        self.nbp, self.Sp = nbp, Sp
        z = lambda *s: torch.zeros(*s, dtype=torch.float64, device=dev)
        self.aug_b = None if self.is_train else z(M, nb, gp.Kp)
        self.vb = z(M, n + 1, nbp)
        self.vb_t = z(M, nb, gp.np_)
        self.lb = z(M, nb, nbp)
        self.lb_inv = z(M, nb, nbp)
        self.lb_inv_t = z(M, nb, nbp)
        self.zb_pad = z(S, nbp)
        self.zb_t = z(nb, Sp)
        self.mean_b = z(M, nbp)
        self.best_f_pad = z(M, Sp)
        self.qmat = z(M, nb, gp.np_) if self.is_train else None
        self.qmat_t = z(M, n, nbp) if self.is_train else None
        S1p = pad8(S + 1)
        self.fwd_cat = z(M, S1p + n, gp.np_) if self.is_train else None
        self.bwd_cat = z(M, n, S1p + gp.np_) if self.is_train else None
        self.info = torch.zeros(M, dtype=torch.int32, device=dev)
        self.c = BoNEI(
            n_b=nb, S=S, baseline_is_train=int(self.is_train), X_b=ptr(self.X_b), Z_b=ptr(self.Z_b), jitter=None,
            aug_b=ptr(self.aug_b), vb=ptr(self.vb), vb_t=ptr(self.vb_t), lb=ptr(self.lb), lb_inv=ptr(self.lb_inv),
            lb_inv_t=ptr(self.lb_inv_t), zb_pad=ptr(self.zb_pad), zb_t=ptr(self.zb_t), mean_b=ptr(self.mean_b),
            best_f=ptr(self.best_f_pad), qmat=ptr(self.qmat), qmat_t=ptr(self.qmat_t),
            fwd_cat=ptr(self.fwd_cat), bwd_cat=ptr(self.bwd_cat), reduction=int(reduction), no_baseline=0,
            obj_weight=float(obj_weight), local_derivs=_local_derivs_default(reduction),
        )  # fmt: skip
        self.jitter = torch.zeros(M, dtype=torch.float64, device=dev)
        if not setup:
            if z_producer is not None:  # allocate-only state (factors arrive by broadcast): just take the base samples
                self.Z_b.copy_(_f64(z_producer(), dev))
            return
        lib = _lib.lib()
        work = _lib.workspace(lib.bo_nei_setup_workspace_bytes(C.byref(gp.c), C.byref(self.c)), dev)

        def run(jitter: Optional[Tensor]) -> Tensor:
            self._jitter = jitter
            self.c.jitter = ptr(jitter)
            check(lib.bo_nei_setup(gp.ctx, C.byref(gp.c), C.byref(self.c), work.data_ptr(), work.numel(),
                                   ptr(self.info), _lib.stream_ptr(dev)))  # fmt: skip
            return self.info.clone()

        drawn: dict = {}
        drawer = None
        if z_producer is not None:  # the host draws the base samples on a thread while this one enqueues and awaits phase 1
            import threading

            def draw() -> None:
                try:
                    drawn["z"] = z_producer()
                except BaseException as e:  # re-raised on the constructing thread
                    drawn["error"] = e

            drawer = threading.Thread(target=draw, daemon=True)
            drawer.start()

        def finish_samples() -> None:  # phase 2: the Z_b-dependent pieces, once the host has produced the base samples
            if drawer is None:
                return
            drawer.join()
            if "error" in drawn:
                raise drawn["error"]
            z = _f64(drawn["z"], dev)
            if z.shape != self.Z_b.shape:
                raise ValueError(f"base samples {tuple(z.shape)} do not match {tuple(self.Z_b.shape)}")
            self.Z_b.copy_(z)
            self.c.setup_phase = 2
            try:
                run(self._jitter)
            finally:
                self.c.setup_phase = 0

        self.c.setup_phase = 0 if z_producer is None else 1
        if forced_jitter is not None:
            # the joint-factorisation fallback (acquisition._FusedImprovementBase._forward_resolved): ONE attempt with this
            # jitter on the baseline block and, in the forward kernel, the same jitter on every q x q block
            self.jitter = torch.full((M,), float(forced_jitter), dtype=torch.float64, device=dev)
            self.setup_failed = bool(run(self.jitter).any())
            self.c.schur_jitter, self.c.schur_fixed = float(forced_jitter), 1
            finish_samples()
            return
        self.jitter = _retry_with_jitter(run, M, dev, "posterior covariance at X_baseline")
        finish_samples()

    def factor_buffers(self):
        """Device buffers bo_nei_setup fills (broadcast payload)."""
        bufs = [self.vb, self.vb_t, self.lb, self.lb_inv, self.lb_inv_t, self.zb_pad, self.zb_t, self.mean_b,
                self.best_f_pad, self.jitter]  # fmt: skip
        bufs += [] if self.aug_b is None else [self.aug_b]
        return bufs + ([] if self.qmat is None else [self.qmat, self.qmat_t, self.fwd_cat, self.bwd_cat])

    @property
    def baseline_L(self) -> Tensor:
        return self.lb[:, :, : self.n_b]

    @property
    def best_f(self) -> Tensor:
        return self.best_f_pad[:, : self.S]

    # -- CUDA-graph replay of the fixed launch sequence (small configs are launch-bound) ---------------
    def graphed(self, B: int, q: int, d: int, Z_q: Tensor, tau_relu: float, tau_max: float, with_grad: bool):
        """Cached ``GraphedEval`` for this shape, or None when graphs are disabled (BOTORCH_B200_GRAPHS=0)."""
        if os.environ.get("BOTORCH_B200_GRAPHS", "1") == "0":
            return None
        key = (B, q, bool(with_grad), Z_q.data_ptr(), float(tau_relu), float(tau_max))
        cache = self.__dict__.setdefault("_graphs", {})
        ge = cache.get(key)
        if ge is None:
            # Capture pays for itself only when a shape recurs MANY times.  Measured (tools/graph_cost.py,
            # profiles/r02_graph_cost.json): the first capture in a process costs 30-46 ms and later ones 2-4 ms, while a
            # replay saves at most ~20 us per evaluation over the eager launches when the caller does not synchronise per
            # round (C1, B = 8: 0.098 -> 0.077 ms) and nothing at C3 shapes -- so the shrinking-batch L-BFGS loop of one
            # acquisition (20-40 rounds per shape) lost 22 ms of 150 to captures it never amortised.  A key is captured at
            # its GRAPH_MIN_USES-th use (long SciPy loops on small problems); one-shot and short-lived shapes stay eager.
            seen = self.__dict__.setdefault("_graph_seen", {})
            seen[key] = seen.get(key, 0) + 1
            min_uses = int(os.environ.get("BOTORCH_B200_GRAPH_AFTER", GRAPH_MIN_USES))
            if seen[key] < min_uses or self.workspace_bytes(B, q, with_grad) > GRAPH_MAX_WORKSPACE_BYTES:
                return None
            if len(cache) >= 6:  # each entry pins its workspace; keep the most recent shapes only
                cache.pop(next(iter(cache)))
            ge = cache[key] = GraphedEval(self, B, q, d, Z_q, tau_relu, tau_max, with_grad)
        return ge

    def workspace_bytes(self, B: int, q: int, with_grad: bool) -> int:
        return _lib.lib().bo_qlognei_workspace_bytes(C.byref(self.gp.c), C.byref(self.c), B, q, int(with_grad))

    def forward(self, X: Tensor, Z_q: Tensor, tau_relu: float, tau_max: float, with_grad: bool,
                work: Optional[Tensor] = None, out: Optional[Tuple[Tensor, Tensor]] = None):  # fmt: skip
        """X: B x q x d, Z_q: S x q -> (acq [M x B], info [M x B] int32, workspace).  ``out``: caller-owned (acq, info)."""
        B, q, _ = X.shape
        dev = self.device
        if out is None:
            acq = torch.empty(self.gp.M, B, dtype=torch.float64, device=dev)
            info = torch.empty(self.gp.M, B, dtype=torch.int32, device=dev)
        else:
            acq, info = out
        if work is None:
            work = _lib.workspace(self.workspace_bytes(B, q, with_grad), dev)
        check(_lib.lib().bo_qlognei_fwd(self.gp.ctx, C.byref(self.gp.c), C.byref(self.c), B, q, ptr(X), ptr(Z_q),
                                        tau_relu, tau_max, ptr(acq), ptr(info), work.data_ptr(), work.numel(),
                                        int(with_grad), _lib.stream_ptr(dev)))  # fmt: skip
        return acq, info, work

    def backward(self, X: Tensor, Z_q: Tensor, tau_relu: float, tau_max: float, grad_acq: Tensor, work: Tensor,
                 out: Optional[Tensor] = None) -> Tensor:  # fmt: skip
        """grad_acq: [M x B] -> grad_X [M x B x q x d] (per model), written into ``out`` when given."""
        B, q, d = X.shape
        grad_X = torch.empty(self.gp.M, B, q, d, dtype=torch.float64, device=self.device) if out is None else out
        check(_lib.lib().bo_qlognei_bwd(self.gp.ctx, C.byref(self.gp.c), C.byref(self.c), B, q, ptr(X), ptr(Z_q),
                                        tau_relu, tau_max, ptr(grad_acq.contiguous()), ptr(grad_X), work.data_ptr(),
                                        work.numel(), _lib.stream_ptr(self.device)))  # fmt: skip
        return grad_X


    # -- outcome constraints across several models: sample rows of the workspace, external sample adjoints --
    def keep_samples(self) -> "NEIState":
        """The caller reads the samples after a forward (outcome constraints, generic objectives): leave them in place."""
        self.c.local_derivs = 0
        return self

    def samples_view(self, work: Tensor, B: int, q: int) -> Tensor:
        """[S x pitch] view of the cached-Cholesky samples a gradient-carrying forward left in ``work`` (M = 1)."""
        if self.c.local_derivs:
            raise _lib.BoError("samples_view needs keep_samples(): this state leaves local derivatives in the sample rows")
        off, pitch = C.c_int64(), C.c_int64()
        check(_lib.lib().bo_qlognei_samples_view(C.byref(self.gp.c), C.byref(self.c), B, q, C.byref(off), C.byref(pitch)))
        flat = work[off.value * 8 : (off.value + self.S * pitch.value) * 8].view(torch.float64)
        return flat.view(self.S, pitch.value)

    def backward_external(self, X: Tensor, Z_q: Tensor, tau_relu: float, tau_max: float, work: Tensor) -> Tensor:
        """Reverse pass with d loss / d f supplied in the sample rows of ``work`` -> grad_X [M x B x q x d]."""
        B, q, d = X.shape
        grad_X = torch.empty(self.gp.M, B, q, d, dtype=torch.float64, device=self.device)
        self.c.external_df = 1
        try:
            check(_lib.lib().bo_qlognei_bwd(self.gp.ctx, C.byref(self.gp.c), C.byref(self.c), B, q, ptr(X), ptr(Z_q),
                                            tau_relu, tau_max, None, ptr(grad_X), work.data_ptr(), work.numel(),
                                            _lib.stream_ptr(self.device)))  # fmt: skip
        finally:
            self.c.external_df = 0
        return grad_X


class EIState(NEIState):
    """qEI / qLogEI: no baseline set, a fixed incumbent ``best_f`` ([M] or scalar) instead of per-sample baseline maxima.

    Same forward / backward entry points; the stacked operands hold only the mean row and the root of
    ys^2 (K + s2 I)^-1 (``bo_ei_setup``), so the q x q posterior covariance is ys^2 k(X,X) - U^T U.
    """

    def __init__(self, gp: GPState, best_f: Tensor, S: int, setup: bool = True, reduction: int = REDUCE_LOG,
                 obj_weight: float = 1.0) -> None:  # fmt: skip
        dev = gp.device
        self.gp, self.device = gp, dev
        self.S, self.n_b, self.is_train = int(S), 0, False
        n, M = gp.n, gp.M
        self.Sp = pad8(self.S)
        z = lambda *s: torch.zeros(*s, dtype=torch.float64, device=dev)
        bf = torch.as_tensor(best_f, dtype=torch.float64, device=dev).reshape(-1)
        if bf.numel() not in (1, M):
            raise ValueError(f"best_f must be a scalar or one value per model ({M}), got {tuple(bf.shape)}")
        self.best_f_pad = bf.expand(M).reshape(M, 1).expand(M, self.Sp).contiguous()
        self.fwd_cat = z(M, 8 + n, gp.np_)
        self.bwd_cat = z(M, n, 8 + gp.np_)
        self.info = torch.zeros(M, dtype=torch.int32, device=dev)
        self.c = BoNEI(n_b=0, S=self.S, baseline_is_train=0, best_f=ptr(self.best_f_pad), fwd_cat=ptr(self.fwd_cat),
                       bwd_cat=ptr(self.bwd_cat), reduction=int(reduction), no_baseline=1,
                       obj_weight=float(obj_weight), local_derivs=_local_derivs_default(reduction))  # fmt: skip
        if not setup:
            return
        lib = _lib.lib()
        work = _lib.workspace(lib.bo_ei_setup_workspace_bytes(C.byref(gp.c), C.byref(self.c)), dev)
        check(lib.bo_ei_setup(gp.ctx, C.byref(gp.c), C.byref(self.c), work.data_ptr(), work.numel(), ptr(self.info),
                              _lib.stream_ptr(dev)))  # fmt: skip
        bad = int(self.info.max())
        if bad != 0:
            raise NotPSDError(f"root of the train precision matrix: pivot {bad} not positive")

    def factor_buffers(self):
        return [self.best_f_pad, self.fwd_cat, self.bwd_cat]


class GraphedEval:
    """bo_qlognei_fwd (+ bo_qlognei_bwd) for one (B, q) shape captured into CUDA graphs over static buffers.

    The library never allocates or synchronises, so its launch sequence captures as-is; replay costs one graph launch
    instead of ~11 kernel launches through ctypes.  ``busy`` marks a forward whose backward has not run yet: a second
    gradient-carrying forward in that window must not reuse the workspace (the caller falls back to the eager path).
    """

    def __init__(self, nei: "NEIState", B: int, q: int, d: int, Z_q: Tensor, tau_relu: float, tau_max: float,
                 with_grad: bool) -> None:  # fmt: skip
        dev = nei.device
        self._nei = nei
        self.X = torch.zeros(B, q, d, dtype=torch.float64, device=dev)
        self.busy = False
        self.with_grad = with_grad
        args = (self.X, Z_q, tau_relu, tau_max)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):  # eager warm-up: one-time function attributes are set outside the capture
            _, _, w = nei.forward(*args, with_grad=with_grad)
            if with_grad:
                nei.backward(*args, torch.zeros(nei.gp.M, B, dtype=torch.float64, device=dev), w)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.fwd = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.fwd):
            self.acq, self.info, self.work = nei.forward(*args, with_grad=with_grad)
            # (min, max) of the failure flags, reduced inside the graph: the host reads two ints per call
            self.info_minmax = torch.stack(torch.aminmax(self.info)).to(torch.int32)
        self.bwd = None
        if with_grad:
            self.grad_in = torch.zeros(nei.gp.M, B, dtype=torch.float64, device=dev)
            self.bwd = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.bwd, pool=self.fwd.pool()):
                self.grad_X = nei.backward(*args, self.grad_in, self.work)

    def run_forward(self, X: Tensor):
        self.X.copy_(X)
        self.fwd.replay()
        self._nei.last_graph_info = self.info  # per-(model, t-batch) flags of this replay (static buffer)
        self.busy = self.with_grad
        return self.acq.clone(), self.info_minmax.detach()  # alias of a static buffer: consumed (copied to the host) before the next replay

    def run_backward(self, grad_acq: Tensor) -> Tensor:
        self.grad_in.copy_(grad_acq)
        self.bwd.replay()
        self.busy = False
        return self.grad_X.sum(dim=0)


# -------------------------------------------------------------------------------------------------
# dense MVN helpers for the posterior API (posteriors/gpytorch.py:85-149: root of Sigma, mu + root z)
# -------------------------------------------------------------------------------------------------
def mvn_root(cov: Tensor, what: str = "posterior covariance") -> Tensor:
    """Lower Cholesky root of a batch of covariance matrices [... x r x r] with psd_safe_cholesky's
    retry schedule; blocked right-looking factorisation inside libbotorch_b200 (bo_potrf)."""
    dev = cov.device
    r = cov.shape[-1]
    batch_shape = cov.shape[:-2]
    nb = max(int(torch.Size(batch_shape).numel()), 1)
    if bool(torch.isnan(cov).any()):
        raise NanError(f"cholesky: {int(torch.isnan(cov).sum())} of {cov.numel()} elements are NaN.")
    rp = pad8(r)
    src = _f64(cov, dev).reshape(nb, r, r)
    lib = _lib.lib()
    work = _lib.workspace(lib.bo_potrf_workspace_bytes(r, nb), dev)
    info = torch.zeros(nb, dtype=torch.int32, device=dev)
    A = torch.zeros(nb, r, rp, dtype=torch.float64, device=dev)

    def run(jitter: Optional[Tensor]) -> Tensor:
        A.zero_()
        A[:, :, :r] = src
        if jitter is not None:
            A[:, :, :r].diagonal(dim1=-2, dim2=-1).add_(jitter.unsqueeze(-1))
        check(lib.bo_potrf(_lib.ctx(dev), r, nb, ptr(A), rp, r * rp, None, work.data_ptr(), work.numel(), ptr(info),
                           _lib.stream_ptr(dev)))  # fmt: skip
        return info.clone()

    _retry_with_jitter(run, nb, dev, what)
    return A[:, :, :r].reshape(*batch_shape, r, r)


def root_apply(L: Tensor, mean: Tensor, Z: Tensor) -> Tensor:
    """samples[s, ..., i] = mean[..., i] + sum_j L[..., i, j] Z[s, ..., j].

    L: [... x r x r] lower, mean: [... x r], Z: [S x ... x r] (batch dims may be broadcast views).
    One triangular DMMA GEMM per batch member (bo_dgemm with BO_TRI_A_LOWER)."""
    dev = L.device
    r = L.shape[-1]
    batch_shape = L.shape[:-2]
    nb = max(int(torch.Size(batch_shape).numel()), 1)
    S = Z.shape[0]
    rp, Sp = pad8(r), pad8(S)
    Lp = torch.zeros(nb, r, rp, dtype=torch.float64, device=dev)
    Lp[:, :, :r] = L.reshape(nb, r, r)
    shared = all(st == 0 or sz == 1 for st, sz in zip(Z.stride()[1:-1], Z.shape[1:-1]))
    if shared:  # one base-sample set for every batch member (t-batch collapsed samplers)
        Zt = torch.zeros(1, r, Sp, dtype=torch.float64, device=dev)
        Zt[0, :, :S] = Z[(slice(None),) + (0,) * (Z.dim() - 2)].to(torch.float64).mT
        strideZ = 0
    else:
        Zt = torch.zeros(nb, r, Sp, dtype=torch.float64, device=dev)
        Zt[:, :, :S] = Z.expand(S, *batch_shape, r).reshape(S, nb, r).permute(1, 2, 0)
        strideZ = r * Sp
    out = torch.empty(nb, r, Sp, dtype=torch.float64, device=dev)
    check(_lib.lib().bo_dgemm(_lib.ctx(dev), r, S, r, nb, 1.0, ptr(Lp), rp, r * rp, ptr(Zt), Sp, strideZ, 0.0, ptr(out),
                              Sp, r * Sp, 1, _lib.stream_ptr(dev)))  # fmt: skip
    samples = out[:, :, :S].permute(2, 0, 1).reshape(S, *batch_shape, r)
    return samples + mean


class _RootSample(torch.autograd.Function):
    """samples = mean + chol(cov) z, differentiable w.r.t. mean and cov (``MultivariateNormal.rsample`` with base samples,
    posteriors/gpytorch.py:121-165).  Forward: blocked Cholesky + triangular GEMM of libbotorch_b200.  Backward: the
    Cholesky adjoint  cov_bar = sym(L^-T Phi(L^T L_bar) L^-1),  L_bar = sum_s samples_bar_s z_s^T  (small r x r algebra)."""

    @staticmethod
    def forward(ctx, mean: Tensor, cov: Tensor, Z: Tensor):
        L = mvn_root(cov.detach())
        ctx.save_for_backward(L, Z)
        return root_apply(L, mean.detach(), Z)

    @staticmethod
    def backward(ctx, g: Tensor):
        L, Z = ctx.saved_tensors
        g = g.to(torch.float64)
        mean_bar = g.sum(dim=0)
        Lbar = torch.einsum("s...i,s...j->...ij", g, Z.to(torch.float64).expand_as(g)).tril()
        P = (L.mT @ Lbar).tril()
        P = P - 0.5 * torch.diag_embed(P.diagonal(dim1=-2, dim2=-1))
        Sb = torch.linalg.solve_triangular(L.mT, P, upper=True)  # L^-T P
        Sb = torch.linalg.solve_triangular(L, Sb, upper=False, left=False)  # (L^-T P) L^-1
        return mean_bar, 0.5 * (Sb + Sb.mT), None


def root_sample(mean: Tensor, cov: Tensor, Z: Tensor) -> Tensor:
    """mean [... x r], cov [... x r x r], Z [S x ... x r] -> samples [S x ... x r]; carries gradients when its inputs do."""
    if torch.is_grad_enabled() and (mean.requires_grad or cov.requires_grad):
        return _RootSample.apply(mean, cov, Z)
    return root_apply(mvn_root(cov.detach()), mean.detach(), Z)


class _PosteriorJoint(torch.autograd.Function):
    """mean/covariance blocks of B t-batches with the hand-written reverse pass (bo_posterior_joint_bwd)."""

    @staticmethod
    def forward(ctx, X: Tensor, st: GPState):
        Xd = _f64(X.detach(), st.device)
        mean, cov = st.posterior_joint(Xd)
        ctx.st = st
        ctx.save_for_backward(Xd)
        ctx.in_dtype = X.dtype
        return mean, cov

    @staticmethod
    def backward(ctx, mean_bar: Tensor, cov_bar: Tensor):
        (Xd,) = ctx.saved_tensors
        st: GPState = ctx.st
        B, q, d = Xd.shape
        lib = _lib.lib()
        zeros = lambda *s: torch.zeros(*s, dtype=torch.float64, device=st.device)
        mean_bar = zeros(st.M, B, q) if mean_bar is None else _f64(mean_bar, st.device)
        cov_bar = zeros(st.M, B, q, q) if cov_bar is None else _f64(cov_bar, st.device)
        grad = torch.empty(st.M, B, q, d, dtype=torch.float64, device=st.device)
        work = _lib.workspace(lib.bo_posterior_bwd_workspace_bytes(C.byref(st.c), B, q), st.device)
        check(lib.bo_posterior_joint_bwd(st.ctx, C.byref(st.c), B, q, ptr(Xd), ptr(mean_bar), ptr(cov_bar), ptr(grad),
                                         work.data_ptr(), work.numel(), _lib.stream_ptr(st.device)))  # fmt: skip
        return grad.sum(dim=0).to(ctx.in_dtype), None


def posterior_joint(st: GPState, X: Tensor):
    """X: B x q x d -> (mean [M x B x q], cov [M x B x q x q]); differentiable w.r.t. X."""
    if X.device != st.device:
        raise _lib.BoError(f"X is on {X.device} but the model factors live on {st.device}; there is no CPU path")
    if torch.is_grad_enabled() and X.requires_grad:
        if X.shape[-2] > MAX_Q:
            raise _lib.BoError(f"the posterior of a {X.shape[-2]}-point joint set is not differentiable w.r.t. X on this "
                               f"path (q <= {MAX_Q}); evaluate it under torch.no_grad()")
        return _PosteriorJoint.apply(X, st)
    return st.posterior_joint(X.detach())
