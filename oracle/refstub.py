"""Load the reference's own importable modules verbatim (container only).

TEST INFRASTRUCTURE.  ``import botorch`` fails on this image because
botorch/__init__.py:7 pulls in gpytorch, which is not installed.  The modules
on the hot path that do not touch gpytorch import fine once the package
``__init__`` files are bypassed with namespace stubs whose ``__path__`` points
at /root/reference/botorch.  Used ONLY by
``tests/golden/make_golden.py`` (fixture generation) and by not-gpu tests that
skip when /root/reference is absent (it does not exist on the GPU box).
"""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("BOTORCH_REFERENCE_ROOT", "/root/reference")

_SUBPACKAGES = [
    "utils",
    "sampling",
    "posteriors",
    "test_functions",
    "optim",
    "optim.utils",
    "acquisition",
    "models",
    "generation",
]


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "botorch"))


def load():
    """Install the stubs and return a namespace of the verbatim modules."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    root = os.path.join(REFERENCE_ROOT, "botorch")
    if "botorch" in sys.modules and not getattr(sys.modules["botorch"], "_is_refstub", False):
        raise RuntimeError("a real `botorch` is already imported; refuse to shadow it")

    def _stub(name: str, path: str) -> None:
        if name in sys.modules:
            return
        m = types.ModuleType(name)
        m.__path__ = [path]
        m._is_refstub = True
        sys.modules[name] = m

    _stub("botorch", root)
    for sub in _SUBPACKAGES:
        _stub("botorch." + sub, os.path.join(root, sub.replace(".", "/")))

    import botorch.exceptions  # noqa: F401  (real package, no gpytorch dependency)
    import botorch.sampling.qmc as qmc
    import botorch.test_functions.synthetic as synthetic
    import botorch.utils.safe_math as safe_math
    import botorch.utils.sampling as sampling
    import botorch.utils.transforms as transforms

    return types.SimpleNamespace(
        safe_math=safe_math,
        qmc=qmc,
        sampling=sampling,
        synthetic=synthetic,
        transforms=transforms,
    )
