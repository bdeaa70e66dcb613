"""Exact-GP models whose ``posterior`` runs on the fused CUDA path.

Drop-in for the reference's model contract on this path:
  * ``Model``                          botorch/models/model.py:60-266 (posterior, batch_shape, num_outputs,
                                        _is_fully_bayesian / _is_ensemble, eval/train)
  * ``SingleTaskGP``                   botorch/models/gp_regression.py:68-254 with the default modules of
                                        botorch/models/utils/gpytorch_modules.py:74-127 (ARD RBF without outputscale,
                                        homoskedastic noise >= 1e-4, ConstantMean) and the default ``Standardize``
                                        outcome transform (botorch/models/transforms/outcome.py:214-447)
  * ``SaasFullyBayesianSingleTaskGP``  botorch/models/fully_bayesian.py:315-546 (M batched hyperparameter samples of
                                        ScaleKernel(Matern-5/2 ARD); MCMC dim at -3; GaussianMixturePosterior)
  * posterior                          botorch/models/gpytorch.py:405-466 under gpt_posterior_settings
                                        (botorch/models/utils/assorted.py:287-297: train caches detached, so
                                        gradients flow to X only)

Hyperparameters are given, loaded from a reference ``state_dict()`` (same keys), or fitted with
``botorch_b200.fit.fit_gpytorch_mll`` (exact MLL value + gradient on the device); NUTS is outside the path.
All arithmetic happens in libbotorch_b200; there is no CPU path.
"""
from __future__ import annotations

import math
import warnings
from typing import Any, Dict, List, Mapping, Optional, Union

import torch
from torch import Tensor
from torch.nn import Module

from . import engine
from .exceptions import (
    BotorchTensorDimensionError,
    InputDataError,
    InputDataWarning,
    NumericalWarning,
    UnsupportedError,
)
from .posteriors import MCMC_DIM, DenseMVN, FusedGPPosterior, GaussianMixturePosterior

MIN_INFERRED_NOISE_LEVEL = 1e-4  # gpytorch_modules.py:29
MIN_LENGTHSCALE = 2.5e-2  # gpytorch_modules.py:122


def default_lengthscale(d: int) -> float:
    """Mode of the dimension-scaled LogNormal(sqrt2 + log(d)/2, sqrt3) prior (gpytorch_modules.py:118-124)."""
    return max(math.exp(math.sqrt(2.0) + 0.5 * math.log(d) - 3.0), MIN_LENGTHSCALE)


class Standardize(Module):
    """Outcome standardisation, single output (``means``/``stdvs``/``_stdvs_sq`` buffers as in the reference)."""

    _is_linear = True

    def __init__(self, m: int = 1, min_stdv: float = 1e-8) -> None:
        super().__init__()
        if m != 1:
            raise UnsupportedError("the fused path models a single outcome (m = 1)")
        self.register_buffer("means", torch.zeros(1, 1))
        self.register_buffer("stdvs", torch.ones(1, 1))
        self.register_buffer("_stdvs_sq", torch.ones(1, 1))
        self.register_buffer("_is_trained", torch.tensor(False))
        self._min_stdv = min_stdv

    def forward(self, Y: Tensor, Yvar: Optional[Tensor] = None):
        if self.training:
            if Y.shape[-2] < 1:
                raise ValueError(f"Can't standardize with no observations. {Y.shape=}.")
            stdvs = Y.std(dim=-2, keepdim=True) if Y.shape[-2] > 1 else torch.ones_like(Y[..., :1, :])
            stdvs = stdvs.where(stdvs >= self._min_stdv, torch.full_like(stdvs, 1.0))
            self.means, self.stdvs, self._stdvs_sq = Y.mean(dim=-2, keepdim=True), stdvs, stdvs.pow(2)
            self._is_trained = torch.tensor(True)
        Y_tf = (Y - self.means) / self.stdvs
        return Y_tf, (None if Yvar is None else Yvar / self._stdvs_sq)

    def untransform(self, Y: Tensor, Yvar: Optional[Tensor] = None):
        return self.means + self.stdvs * Y, (None if Yvar is None else Yvar * self._stdvs_sq)


class _Holder(Module):
    """Parameter container so that ``state_dict()`` keys read like the reference's module tree."""


class Model(Module):
    """The reference's abstract model interface (only what callers on the path use)."""

    _is_fully_bayesian = False
    _is_ensemble = False

    def posterior(self, X, output_indices=None, observation_noise=False, posterior_transform=None):  # pragma: no cover
        raise NotImplementedError

    @property
    def batch_shape(self) -> torch.Size:  # pragma: no cover
        raise NotImplementedError

    @property
    def num_outputs(self) -> int:
        return 1

    def transform_inputs(self, X: Tensor, input_transform: Optional[Module] = None) -> Tensor:
        return X


def _validate_tensor_args(X: Tensor, Y: Tensor) -> None:
    if X.dim() != Y.dim():
        if X.dim() - Y.dim() == 1 and X.shape[:-1] == Y.shape:
            raise BotorchTensorDimensionError(
                f"An explicit output dimension is required for targets. Expected Y with dimension {X.dim()} "
                f"(got {Y.dim()=})."
            )
        raise BotorchTensorDimensionError(
            f"Expected X and Y to have the same number of dimensions (got X with dimension {X.dim()} and Y with "
            f"dimension {Y.dim()})."
        )
    if X.dtype != Y.dtype:
        raise InputDataError(f"Expected all inputs to share the same dtype. Got {X.dtype} for X, {Y.dtype} for Y.")
    if X.dtype != torch.float64:
        warnings.warn(
            f"The model inputs are of type {X.dtype}. It is strongly recommended to use double precision in BoTorch, "
            "as this improves both precision and stability and can help avoid numerical errors.",
            InputDataWarning,
            stacklevel=3,
        )


class _FusedExactGP(Model):
    """Shared machinery: training data, lazily built device factors, the posterior call."""

    _kernel = "rbf"

    def _init_data(self, train_X: Tensor, train_Y: Tensor, outcome_transform, train_Yvar: Optional[Tensor] = None) -> None:
        _validate_tensor_args(train_X, train_Y)
        if train_X.dim() != 2 or train_Y.shape[-1] != 1:
            raise UnsupportedError("the fused path takes unbatched n x d inputs and n x 1 targets")
        if train_Yvar is not None and train_Yvar.shape != train_Y.shape:
            raise BotorchTensorDimensionError(
                f"An explicit output dimension is required for observation noise. Expected Yvar with shape: "
                f"{tuple(train_Y.shape)} (got {tuple(train_Yvar.shape)})."
            )
        if outcome_transform is not None:
            outcome_transform.train()
            train_Y, train_Yvar = outcome_transform(train_Y, train_Yvar)  # Yvar / stdvs^2 (outcome.py:289-307)
            outcome_transform.eval()
            self.outcome_transform = outcome_transform
        self.train_inputs = (train_X,)
        self.train_targets = train_Y.squeeze(-1)
        # fixed per-observation noise (FixedNoiseGaussianLikelihood, gp_regression.py:186-195), model space
        self.train_Yvar = None
        if train_Yvar is not None:  # gpytorch FixedGaussianNoise: settings.min_fixed_noise (1e-6 in double, 1e-4 in float)
            floor = 1e-6 if train_Yvar.dtype == torch.float64 else 1e-4
            if bool((train_Yvar < floor).any()):
                warnings.warn(
                    "Very small noise values detected. This will likely lead to numerical instabilities. Rounding "
                    f"small noise values up to {floor}.",
                    NumericalWarning,
                    stacklevel=3,
                )
            self.train_Yvar = train_Yvar.detach().squeeze(-1).clamp_min(floor)
        self._state: Optional[engine.GPState] = None

    # -- hyperparameters (subclasses) -----------------------------------------------------------
    def _hypers(self) -> Dict[str, Optional[Tensor]]:  # pragma: no cover
        raise NotImplementedError

    def train(self, mode: bool = True):
        if mode:
            self._state = None  # hyperparameters may change: drop the device factors
        return super().train(mode)

    # -- device factors (the prediction caches of gpytorch's DefaultPredictionStrategy) ----------
    def _build_state(self, setup: bool = True) -> engine.GPState:
        """Allocate the device factors; ``setup=False`` skips the factorisation (a peer rank broadcasts them)."""
        h = self._hypers()
        ot = getattr(self, "outcome_transform", None)
        return engine.GPState(
            self.train_inputs[0].detach(),
            self.train_targets.detach(),
            self._kernel,
            h["lengthscale"],
            h["noise"],
            h["mean"],
            h.get("outputscale"),
            y_mean=float(ot.means) if ot is not None else 0.0,
            y_std=float(ot.stdvs) if ot is not None else 1.0,
            setup=setup,
            noise_vec=getattr(self, "train_Yvar", None),
        )

    @property
    def gp_state(self) -> engine.GPState:
        if self._state is None:
            self._state = self._build_state()
        return self._state

    def _posterior_blocks(self, X: Tensor, observation_noise: Union[bool, Tensor]):
        d = self.train_inputs[0].shape[-1]
        if X.dim() < 2 or X.shape[-1] != d:
            raise BotorchTensorDimensionError(f"Expected X of shape (batch_shape) x q x {d}, got {tuple(X.shape)}.")
        self.eval()
        st = self.gp_state
        batch_shape, q = X.shape[:-2], X.shape[-2]
        Xf = X.reshape(-1, q, d)
        mean, cov = engine.posterior_joint(st, Xf)  # [M x B x q], [M x B x q x q]; differentiable w.r.t. X
        if isinstance(observation_noise, Tensor):
            cov = cov + torch.diag_embed(observation_noise.reshape(1, -1, q).to(cov))
        elif observation_noise:
            ys2 = st.y_std**2
            cov = cov + torch.diag_embed((ys2 * st.noise).reshape(-1, 1, 1).expand(st.M, Xf.shape[0], q))
        return mean, cov, batch_shape, q


class SingleTaskGP(_FusedExactGP):
    """Single-outcome exact GP, ARD RBF kernel without outputscale, inferred homoskedastic noise.

    ``lengthscale`` (d), ``noise`` and ``mean_constant`` are the fitted hyperparameters; they default to the
    initial values the reference's module factories use (prior modes).  A ``state_dict()`` of a reference
    ``SingleTaskGP`` loads directly (``covar_module.raw_lengthscale``, ``likelihood.noise_covar.raw_noise``,
    ``mean_module.raw_constant``, ``outcome_transform.*``; the reference's constraints use ``transform=None`` so
    raw values are the values).
    """

    def __new__(cls, train_X=None, train_Y=None, *args, **kwargs):
        if cls is SingleTaskGP and isinstance(train_Y, Tensor) and train_Y.dim() == 2 and train_Y.shape[-1] > 1:
            from .constrained import ModelListGP  # n x m targets: m independent outputs (models/gpytorch.py:405-466)

            return ModelListGP.from_batched(train_X, train_Y, *args, **kwargs)
        return super().__new__(cls)

    def __init__(
        self,
        train_X: Tensor,
        train_Y: Tensor,
        train_Yvar: Optional[Tensor] = None,
        likelihood=None,
        covar_module=None,
        mean_module=None,
        outcome_transform: Union[str, Module, None] = "DEFAULT",
        input_transform=None,
        *,
        lengthscale: Optional[Tensor] = None,
        noise: Union[Tensor, float, None] = None,
        mean_constant: Union[Tensor, float] = 0.0,
        outputscale: Union[Tensor, float, None] = None,
        kernel: str = "rbf",
    ) -> None:
        super().__init__()
        for name, arg in (("likelihood", likelihood), ("covar_module", covar_module),
                          ("mean_module", mean_module), ("input_transform", input_transform)):  # fmt: skip
            if arg is not None:
                raise UnsupportedError(f"`{name}` is outside the fused path (SURVEY.md section 8): pass None")
        if kernel not in engine.KERNEL_IDS:
            raise UnsupportedError(f"kernel must be one of {sorted(engine.KERNEL_IDS)}")
        if isinstance(outcome_transform, str):
            outcome_transform = Standardize(m=1)
        self._kernel = kernel
        if train_Yvar is not None and noise is not None:
            raise ValueError("pass either `train_Yvar` (fixed per-observation noise) or `noise` (inferred homoskedastic level)")
        self._init_data(train_X, train_Y, outcome_transform, train_Yvar)
        d = train_X.shape[-1]
        tk = {"dtype": train_X.dtype, "device": train_X.device}
        ls = torch.full((d,), default_lengthscale(d), **tk) if lengthscale is None else torch.as_tensor(lengthscale, **tk)
        nz = math.exp(-4.0 - 1.0) if noise is None else noise  # mode of LogNormal(-4, 1)
        if self.train_Yvar is not None:  # observation_noise=True adds the mean training noise (models/gpytorch.py:393-400)
            nz = float(self.train_Yvar.mean())
        self.covar_module = _Holder()
        self.covar_module.raw_lengthscale = torch.nn.Parameter(ls.reshape(1, d).clone())
        if outputscale is not None:
            self.covar_module.raw_outputscale = torch.nn.Parameter(torch.as_tensor(outputscale, **tk).reshape(()))
        self.likelihood = _Holder()
        self.likelihood.noise_covar = _Holder()
        self.likelihood.noise_covar.raw_noise = torch.nn.Parameter(torch.as_tensor(nz, **tk).reshape(1).clone())
        self.mean_module = _Holder()
        self.mean_module.raw_constant = torch.nn.Parameter(torch.as_tensor(mean_constant, **tk).reshape(()).clone())

    @property
    def batch_shape(self) -> torch.Size:
        return torch.Size()

    def _hypers(self):
        os = getattr(self.covar_module, "raw_outputscale", None)
        return {
            "lengthscale": self.covar_module.raw_lengthscale.detach().reshape(-1).clamp_min(MIN_LENGTHSCALE),
            "noise": self.likelihood.noise_covar.raw_noise.detach().reshape(()).clamp_min(MIN_INFERRED_NOISE_LEVEL),
            "mean": self.mean_module.raw_constant.detach().reshape(()),
            "outputscale": None if os is None else os.detach().reshape(()),
        }

    def load_state_dict(self, state_dict: Mapping[str, Any], strict: bool = True):
        own = self.state_dict()
        picked = {k: v for k, v in state_dict.items() if k in own}
        for k, v in picked.items():  # buffers of Standardize change shape when untrained -> assign
            if v.shape != own[k].shape:
                picked[k] = v.reshape(own[k].shape)
        missing = [k for k in own if k not in picked]
        if strict and missing:
            raise RuntimeError(f"Missing key(s) in state_dict: {missing}")
        out = super().load_state_dict(picked, strict=False)
        self._state = None
        return out

    def posterior(
        self,
        X: Tensor,
        output_indices: Optional[List[int]] = None,
        observation_noise: Union[bool, Tensor] = False,
        posterior_transform=None,
    ) -> FusedGPPosterior:
        if output_indices not in (None, [0]):
            raise UnsupportedError("single-output model: output_indices must be None")
        mean, cov, batch_shape, q = self._posterior_blocks(X, observation_noise)
        mvn = DenseMVN(mean[0].reshape(*batch_shape, q), cov[0].reshape(*batch_shape, q, q))
        post = FusedGPPosterior(mvn)
        return post if posterior_transform is None else posterior_transform(post)


class SaasFullyBayesianSingleTaskGP(_FusedExactGP):
    """Fully Bayesian single-task GP: M hyperparameter samples evaluated as one batched model.

    ``load_mcmc_samples`` takes the dict the reference's ``SaasPyroModel.postprocess_mcmc_samples`` produces:
    ``mean`` [M], ``lengthscale`` [M x d], ``outputscale`` [M], ``noise`` [M].
    """

    _is_fully_bayesian = True
    _is_ensemble = True
    _kernel = "matern52"

    def __init__(self, train_X: Tensor, train_Y: Tensor, train_Yvar: Optional[Tensor] = None, outcome_transform=None,
                 input_transform=None, pyro_model=None) -> None:  # fmt: skip
        super().__init__()
        if train_Yvar is not None or input_transform is not None or pyro_model is not None:
            raise UnsupportedError("train_Yvar / input_transform / pyro_model are outside the fused path")
        if not (train_X.ndim == train_Y.ndim == 2 and len(train_X) == len(train_Y) and train_Y.shape[-1] == 1):
            raise ValueError("Expected train_X to have shape n x d and train_Y to have shape n x 1")
        self._init_data(train_X, train_Y, outcome_transform)
        self.mean_module = self.covar_module = self.likelihood = None
        self._mcmc: Optional[Dict[str, Tensor]] = None

    def _check_if_fitted(self) -> None:
        if self._mcmc is None:
            raise RuntimeError(
                "Model has not been fitted. You need to call `fit_fully_bayesian_model_nuts` to fit the model."
            )

    def load_mcmc_samples(self, mcmc_samples: Dict[str, Tensor]) -> None:
        X = self.train_inputs[0]
        tk = {"dtype": X.dtype, "device": X.device}
        M = len(mcmc_samples["mean"])
        self._mcmc = {
            "mean": torch.as_tensor(mcmc_samples["mean"], **tk).reshape(M).detach().clone(),
            "lengthscale": torch.as_tensor(mcmc_samples["lengthscale"], **tk).reshape(M, X.shape[-1]).detach().clone(),
            "outputscale": torch.as_tensor(mcmc_samples["outputscale"], **tk).reshape(M).detach().clone(),
            "noise": torch.as_tensor(mcmc_samples["noise"], **tk).reshape(M).clamp_min(MIN_INFERRED_NOISE_LEVEL).detach().clone(),
        }  # fmt: skip
        self._state = None
        self.eval()

    def train(self, mode: bool = True):
        if mode:
            self._mcmc = None  # the reference drops its modules in train mode (fully_bayesian.py:446-452)
        return super().train(mode)

    def _hypers(self):
        self._check_if_fitted()
        return self._mcmc

    @property
    def num_mcmc_samples(self) -> int:
        self._check_if_fitted()
        return len(self._mcmc["outputscale"])

    @property
    def median_lengthscale(self) -> Tensor:
        self._check_if_fitted()
        return self._mcmc["lengthscale"].median(0).values

    @property
    def batch_shape(self) -> torch.Size:
        return torch.Size([self.num_mcmc_samples])

    # reference state-dict keys (test/models/test_fully_bayesian.py:76-89); gpytorch's default Positive /
    # GreaterThan constraints use softplus: value = lower + softplus(raw)
    def state_dict(self, *args, **kwargs):
        sd = super().state_dict(*args, **kwargs)
        if self._mcmc is not None:
            inv = lambda v: v + torch.log(-torch.expm1(-v))  # softplus^-1
            m = self._mcmc
            M, d = m["lengthscale"].shape
            sd["mean_module.raw_constant"] = m["mean"].clone()
            sd["covar_module.raw_outputscale"] = inv(m["outputscale"])
            sd["covar_module.base_kernel.raw_lengthscale"] = inv(m["lengthscale"]).reshape(M, 1, d)
            sd["likelihood.noise_covar.raw_noise"] = inv(m["noise"] - MIN_INFERRED_NOISE_LEVEL).reshape(M, 1)
        return sd

    def load_state_dict(self, state_dict: Mapping[str, Any], strict: bool = True):
        sp = torch.nn.functional.softplus
        M = len(state_dict["mean_module.raw_constant"])
        self.load_mcmc_samples(
            {
                "mean": state_dict["mean_module.raw_constant"],
                "outputscale": sp(state_dict["covar_module.raw_outputscale"]),
                "lengthscale": sp(state_dict["covar_module.base_kernel.raw_lengthscale"]).reshape(M, -1),
                "noise": MIN_INFERRED_NOISE_LEVEL + sp(state_dict["likelihood.noise_covar.raw_noise"]).reshape(M),
            }
        )
        rest = {k: v for k, v in state_dict.items() if k.startswith("outcome_transform.")}
        return super().load_state_dict(rest, strict=False)

    def posterior(
        self,
        X: Tensor,
        output_indices: Optional[List[int]] = None,
        observation_noise: Union[bool, Tensor] = False,
        posterior_transform=None,
    ) -> GaussianMixturePosterior:
        self._check_if_fitted()
        if output_indices not in (None, [0]):
            raise UnsupportedError("single-output model: output_indices must be None")
        mean, cov, batch_shape, q = self._posterior_blocks(X, observation_noise)
        M = mean.shape[0]
        # MCMC dim at -3 (models/fully_bayesian.py:539): batch_shape x M x q
        mean = mean.reshape(M, *batch_shape, q).movedim(0, -2)
        cov = cov.reshape(M, *batch_shape, q, q).movedim(0, MCMC_DIM)
        post = GaussianMixturePosterior(DenseMVN(mean, cov))
        return post if posterior_transform is None else posterior_transform(post)
