"""botorch_b200: the batched GP-posterior + qLogNEI hot path of BoTorch as hand-written sm_100a CUDA.

Only the path lives here: model posterior, base samples, the fused qLogNEI forward/backward,
the multi-start optimiser seam and its multi-GPU sharding.  Python is the host language because the reference is
Python; every number is produced by ``csrc/libbotorch_b200.so`` through the C ABI in ``include/botorch_b200.h``.
There is no CPU fallback: importing is cheap, the first compute call raises if the library or a B200 is missing.
"""
from .acquisition import (  # noqa: F401
    AcquisitionFunction,
    GenericMCObjective,
    IdentityMCObjective,
    LinearMCObjective,
    prune_inferior_points,
    qExpectedImprovement,
    qLogExpectedImprovement,
    qLogNoisyExpectedImprovement,
    qNoisyExpectedImprovement,
)
from .constrained import ModelListGP  # noqa: F401
from .fit import ExactMarginalLogLikelihood, SumMarginalLogLikelihood, fit_gpytorch_mll, fit_gpytorch_mll_scipy  # noqa: F401
from .fully_bayesian import SaasPotential, fit_fully_bayesian_model_nuts, nuts_sample  # noqa: F401
from .generation import MatheronPath, MaxPosteriorSampling, PathwiseThompsonSampling, draw_matheron_paths  # noqa: F401
from .models import SaasFullyBayesianSingleTaskGP, SingleTaskGP, Standardize  # noqa: F401
from .optim import (  # noqa: F401
    gen_batch_initial_conditions,
    gen_candidates_lbfgs,
    gen_candidates_scipy,
    gen_candidates_torch,
    get_best_candidates,
    initialize_q_batch,
    optimize_acqf,
)
from .posteriors import FusedGPPosterior, GaussianMixturePosterior  # noqa: F401
from .sampling import (  # noqa: F401
    IIDNormalSampler,
    SobolQMCNormalSampler,
    draw_sobol_normal_samples,
    draw_sobol_samples,
    get_sampler,
)

__version__ = "0.1.0"


def _register_with_reference_bases() -> None:
    """When the real BoTorch is importable, make `isinstance(x, botorch.<Base>)` hold for the classes of this package
     without inheriting behaviour: the reference's bases are ABCs, so `Base.register(cls)` declares a
    virtual subclass.  `exceptions.py` re-uses the real exception types outright.  No-op when `botorch` is absent (this
    image: gpytorch / linear_operator are not installed)."""
    try:
        from botorch.acquisition.acquisition import AcquisitionFunction as _RefAcq
        from botorch.models.model import Model as _RefModel
        from botorch.posteriors.posterior import Posterior as _RefPosterior
        from botorch.sampling.base import MCSampler as _RefSampler
    except Exception:  # ImportError, or a broken partial install
        return
    from .acquisition import AcquisitionFunction as _Acq
    from .models import Model as _Model
    from .posteriors import Posterior as _Posterior
    from .sampling import MCSampler as _Sampler

    for ref, ours in ((_RefAcq, _Acq), (_RefModel, _Model), (_RefPosterior, _Posterior), (_RefSampler, _Sampler)):
        register = getattr(ref, "register", None)
        if register is not None:
            register(ours)


_register_with_reference_bases()
