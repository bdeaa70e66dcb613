"""Error and warning classes the boundary raises.

Same names as /root/reference/botorch/exceptions/{errors,warnings}.py and, for the
numerical failures that originate in linear_operator / gpytorch in the reference,
linear_operator.utils.errors.{NanError,NotPSDError} and
gpytorch.utils.warnings.NumericalWarning.  If the real `botorch` is importable its
classes are re-used so that `except botorch.exceptions.X` keeps working.
"""
from __future__ import annotations

try:  # pragma: no cover - botorch is not installed on the build image
    from botorch.exceptions.errors import (  # type: ignore
        BotorchError,
        BotorchTensorDimensionError,
        CandidateGenerationError,
        InputDataError,
        OptimizationTimeoutError,
        UnsupportedError,
    )
    from botorch.exceptions.warnings import (  # type: ignore
        BadInitialCandidatesWarning,
        BotorchWarning,
        InputDataWarning,
        NumericsWarning,
        OptimizationWarning,
        SamplingWarning,
    )
except Exception:

    class BotorchError(Exception):
        """Base botorch exception."""

    class CandidateGenerationError(BotorchError):
        pass

    class InputDataError(BotorchError):
        pass

    class UnsupportedError(BotorchError):
        pass

    class BotorchTensorDimensionError(BotorchError):
        pass

    class OptimizationTimeoutError(BotorchError):
        def __init__(self, *args, current_x=None, runtime: float = 0.0, **kwargs):
            super().__init__(*args, **kwargs)
            self.current_x = current_x
            self.runtime = runtime

    class BotorchWarning(Warning):
        pass

    class BadInitialCandidatesWarning(BotorchWarning):
        pass

    class InputDataWarning(BotorchWarning):
        pass

    class OptimizationWarning(BotorchWarning):
        pass

    class SamplingWarning(BotorchWarning):
        pass

    class NumericsWarning(BotorchWarning):
        pass


try:  # pragma: no cover
    from linear_operator.utils.errors import NanError, NotPSDError  # type: ignore
    from linear_operator.utils.warnings import NumericalWarning  # type: ignore
except Exception:

    class NanError(RuntimeError):
        pass

    class NotPSDError(RuntimeError):
        pass

    class NumericalWarning(RuntimeWarning):
        pass
