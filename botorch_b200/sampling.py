"""Base-sample generation for the fused path (host side).

Mirrors the reference's sampler contract so that base samples are bit-identical to BoTorch's:
  * ``NormalQMCEngine``            botorch/sampling/qmc.py:25-101
  * ``draw_sobol_samples`` /
    ``draw_sobol_normal_samples``  botorch/utils/sampling.py:64-136
  * ``manual_seed``                botorch/utils/sampling.py:40-61
  * ``MCSampler``                  botorch/sampling/base.py:25-148
  * ``NormalMCSampler`` & co       botorch/sampling/normal.py:25-209
  * ``get_sampler`` (Sobol when the joint dimension fits SobolEngine.MAXDIM, else iid)
                                    botorch/sampling/get_sampler.py

The Sobol bits come from torch's own ``SobolEngine`` on the host, exactly as in the reference; the
draw happens once per acquisition object and the result is copied to HBM, where the kernels stream
it.  Applying the base samples (f = mu + L z) happens inside the CUDA kernels.
"""
from __future__ import annotations

import math
import os
from contextlib import contextmanager
from typing import Optional, Tuple

import torch
from torch import Tensor
from torch.nn import Module
from torch.quasirandom import SobolEngine

from .exceptions import UnsupportedError


@contextmanager
def manual_seed(seed: Optional[int] = None):
    """Seed torch's global RNG inside the block and restore its state afterwards."""
    saved = None
    if seed is not None:
        saved = torch.random.get_rng_state()
        torch.manual_seed(seed)
    try:
        yield
    finally:
        if saved is not None:
            torch.random.set_rng_state(saved)


def _usable_cpus() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover - non-Linux
        return os.cpu_count() or 1


_SOBOL_THREADS = max(1, min(8, _usable_cpus()))
_sobol_pool = None


def _pool():
    global _sobol_pool
    if _sobol_pool is None:
        from concurrent.futures import ThreadPoolExecutor

        _sobol_pool = ThreadPoolExecutor(max_workers=_SOBOL_THREADS, thread_name_prefix="bo-sobol")
    return _sobol_pool


_state_cache: dict = {}


def _initial_state(dimension: int) -> Tensor:
    """The unscrambled direction-number table of ``SobolEngine(dimension)`` (a pure function of the dimension, read-only
    here: the workers copy the rows they scramble).  One acquisition builds three engines and a BO loop reuses the same
    few dimensions, so the last few tables are kept."""
    state = _state_cache.get(dimension)
    if state is None:
        state = torch.zeros(dimension, SobolEngine.MAXBIT, device=torch.device("cpu"), dtype=torch.long)
        torch._sobol_engine_initialize_state_(state, dimension)
        if len(_state_cache) >= 8:
            _state_cache.pop(next(iter(_state_cache)))
        _state_cache[dimension] = state
    return state


def sobol_draw(dimension: int, n: int, seed: Optional[int] = None, dtype: Optional[torch.dtype] = None,
               columns: Optional[Tensor] = None, first: int = 0, post=None) -> Tensor:  # fmt: skip
    """Points ``first .. first + n - 1`` of ``SobolEngine(dimension, scramble=True, seed=seed)`` -- bit for bit -- or only
    their ``columns`` (a sorted 1-D index tensor), computed without scrambling the dimensions nobody reads.  ``first`` > 0
    is the engine's ``fast_forward(first)`` (integer recurrence only): a rank that evaluates rows [lo, hi) of the raw
    samples draws those rows and no others.

    The reference draws its base samples and raw candidates from torch's engine (sampling/qmc.py:39-105,
    utils/sampling.py:97-140) and so do we; what changes is how the host work is laid out.  Owen scrambling is
    independent per dimension: the random bits are one stream (``shift`` then the lower-triangular matrices, both drawn
    for ALL dimensions so the stream position of every dimension is unchanged), but the 30 x 30 bit-matrix products and
    the Gray-code recurrence touch one dimension at a time.  So (1) the appended-points draw of the cached-Cholesky path
    needs q of the n_b + q dimensions (the baseline columns are overwritten, cached_cholesky.py:167-186) and pays for q;
    (2) a full draw is split by dimension over host threads (torch's ops release the GIL).  At n_b = 2000, q = 8 the
    first goes from ~70 ms to ~10 ms and the second from ~70 ms to ~25 ms on 8 cores; tests/test_host_logic.py holds
    both to ``torch.equal`` with the stock engine.  ``post(block, sel)`` (optional) is an elementwise map applied by each
    worker to its block of columns ``sel`` before it is stored -- the inverse normal CDF, the affine map into a box -- so
    that this work is threaded too (under torchrun ``OMP_NUM_THREADS=1`` makes torch's own elementwise loops serial: the
    42 MB affine map of the C3 raw samples alone took ~30 ms on one core).
    """
    if dimension > SobolEngine.MAXDIM or dimension < 1:
        raise ValueError(f"Supported range of dimensionality for SobolEngine is [1, {SobolEngine.MAXDIM}]")
    dtype = torch.get_default_dtype() if dtype is None else dtype
    bits = SobolEngine.MAXBIT
    cpu = torch.device("cpu")
    state = _initial_state(dimension)
    g = None
    if seed is not None:
        g = torch.Generator()
        g.manual_seed(seed)
    shift_bits = torch.randint(2, (dimension, bits), device=cpu, generator=g)
    ltm = torch.randint(2, (dimension, bits, bits), device=cpu, generator=g)
    pow2 = torch.pow(2, torch.arange(0, bits, device=cpu))
    if columns is None:
        width = dimension
    else:
        columns = columns.to(device=cpu, dtype=torch.long)
        width = columns.numel()
    out = torch.empty(n, width, dtype=dtype)
    if width == 0 or n == 0:
        return out

    def work(a: int, b: int) -> None:
        sel = slice(a, b) if columns is None else columns[a:b]
        st = state[sel].clone(memory_format=torch.contiguous_format)  # the table is shared (cached): scramble a copy
        lower = ltm[sel].tril().contiguous()
        torch._sobol_engine_scramble_(st, lower, b - a)
        quasi = torch.mv(shift_bits[sel], pow2).clone(memory_format=torch.contiguous_format)
        if first == 0:  # point 0 is the shift itself (quasirandom.py:77-113)
            blk = (quasi / 2**bits).to(dtype).reshape(1, -1)
            if n > 1:
                rest, _ = torch._sobol_engine_draw(quasi, n - 1, st, b - a, 0, dtype=dtype)
                blk = torch.cat((blk, rest), dim=0)
        else:
            if first > 1:
                torch._sobol_engine_ff_(quasi, first - 1, st, b - a, 0)
            blk, _ = torch._sobol_engine_draw(quasi, n, st, b - a, first - 1, dtype=dtype)
        out[:, a:b] = blk if post is None else post(blk, sel)

    chunks = max(1, min(_SOBOL_THREADS, width // 16))
    edges = [(width * i) // chunks for i in range(chunks + 1)]
    if chunks == 1:
        work(0, width)
    else:
        for f in [_pool().submit(work, edges[i], edges[i + 1]) for i in range(chunks)]:
            f.result()
    return out


def _normal_from_uniform(u: Tensor) -> Tensor:
    """Inverse normal CDF after shrinking the uniforms away from {0, 1} by one ulp (sampling/qmc.py:92-96)."""
    v = 0.5 + (1 - torch.finfo(u.dtype).eps) * (u - 0.5)
    return torch.erfinv(2 * v - 1) * math.sqrt(2)


class NormalQMCEngine:
    """Quasi-random N(0, I_d) draws from a scrambled Sobol sequence.

    ``inv_transform=True`` maps uniforms through the inverse normal CDF after shrinking them away
    from {0, 1} by one ulp; otherwise consecutive pairs go through Box-Muller.
    """

    def __init__(self, d: int, seed: Optional[int] = None, inv_transform: bool = False) -> None:
        self._d = d
        self._seed = seed
        self._inv_transform = inv_transform
        sobol_dim = d if inv_transform else 2 * math.ceil(d / 2)
        self._sobol_dim = sobol_dim
        self._sobol_engine = None  # torch's engine, built only for a second draw (the first goes through sobol_draw)
        self._drawn = 0

    def draw(self, n: int = 1, out: Optional[Tensor] = None, dtype: Optional[torch.dtype] = None) -> Optional[Tensor]:
        dtype = torch.get_default_dtype() if dtype is None else dtype
        if self._drawn == 0 and self._seed is not None:
            u = sobol_draw(self._sobol_dim, n, seed=self._seed, dtype=dtype)
        else:
            if self._sobol_engine is None:
                self._sobol_engine = SobolEngine(dimension=self._sobol_dim, scramble=True, seed=self._seed)
                if self._drawn:
                    self._sobol_engine.fast_forward(self._drawn)
            u = self._sobol_engine.draw(n, dtype=dtype)
        self._drawn += n
        if self._inv_transform:
            z = _normal_from_uniform(u)
        else:
            even = torch.arange(0, u.shape[-1], 2)
            radius = torch.sqrt(-2 * torch.log(u[:, even]))
            angle = 2 * math.pi * u[:, 1 + even]
            z = (radius.unsqueeze(-1) * torch.stack([torch.cos(angle), torch.sin(angle)], -1)).reshape(n, -1)
            z = z[:, : self._d]
        if out is None:
            return z
        out.copy_(z)
        return None


def draw_sobol_samples(
    bounds: Tensor,
    n: int,
    q: int,
    batch_shape: Optional[torch.Size] = None,
    seed: Optional[int] = None,
    rows: Optional[Tuple[int, int]] = None,
) -> Tensor:
    """``n x batch_shape x q x d`` scrambled-Sobol points inside the box ``bounds`` (2 x d)."""
    batch_shape = torch.Size(batch_shape or ())
    d = bounds.shape[-1]
    lower, width = bounds[0], bounds[1] - bounds[0]
    lo_flat, w_flat = lower.cpu().repeat(q), width.cpu().repeat(q)

    def into_box(u: Tensor, sel) -> Tensor:  # lower + width * u on one block of columns (same two roundings as the reference)
        return lo_flat[sel] + w_flat[sel] * u

    if rows is not None:  # rows [lo, hi) of the n points only (no batch shape): one rank's block of the raw samples
        lo, hi = rows
        x = sobol_draw(q * d, hi - lo, seed=seed, dtype=lower.dtype, first=lo, post=into_box)
        return x.view(hi - lo, q, d).to(device=lower.device)
    x = sobol_draw(q * d, batch_shape.numel() * n, seed=seed, dtype=lower.dtype, post=into_box)
    x = x.view(*batch_shape, n, q, d).to(device=lower.device)
    if len(batch_shape) > 0:  # the sample index leads
        x = x.movedim(len(batch_shape), 0)
    return x


def draw_sobol_normal_samples(
    d: int,
    n: int,
    device: Optional[torch.device] = None,
    dtype: Optional[torch.dtype] = None,
    seed: Optional[int] = None,
) -> Tensor:
    """``n x d`` quasi-random standard normal draws (inverse-CDF Sobol)."""
    return sobol_draw(d, n, seed=seed, dtype=dtype, post=lambda u, sel: _normal_from_uniform(u)).to(device=device)


class MCSampler(Module):
    """Holds the sample shape, the seed and the (t-batch collapsed) ``base_samples`` buffer."""

    def __init__(self, sample_shape: torch.Size, seed: Optional[int] = None) -> None:
        super().__init__()
        if not isinstance(sample_shape, torch.Size):
            raise ValueError(f"Expected `sample_shape` to be a `torch.Size` object, got {sample_shape}.")
        self.sample_shape = sample_shape
        self.seed = seed if seed is not None else int(torch.randint(0, 1000000, (1,)).item())
        self.register_buffer("base_samples", None)

    def forward(self, posterior) -> Tensor:  # pragma: no cover - abstract
        raise NotImplementedError

    def _get_batch_range(self, posterior) -> Tuple[int, int]:
        return getattr(self, "batch_range_override", None) or posterior.batch_range

    def _get_collapsed_shape(self, posterior) -> torch.Size:
        """base-sample shape with every t-batch dimension set to 1 (one draw shared by all t-batches)."""
        full = posterior.base_sample_shape
        lo, hi = self._get_batch_range(posterior)
        collapsed = full[:lo] + torch.Size([1] * len(full[lo:hi])) + full[hi:]
        return self.sample_shape + collapsed

    def _get_extended_base_sample_shape(self, posterior) -> torch.Size:
        return self.sample_shape + posterior.base_sample_shape

    def _instance_check(self, base_sampler) -> None:
        if not isinstance(base_sampler, self.__class__):
            raise RuntimeError(f"Expected `base_sampler` to be an instance of {self.__class__.__name__}.")


class NormalMCSampler(MCSampler):
    """Samplers whose base samples are standard normal."""

    def forward(self, posterior) -> Tensor:
        self._construct_base_samples(posterior=posterior)
        return posterior.rsample_from_base_samples(
            sample_shape=self.sample_shape,
            base_samples=self.base_samples.expand(self._get_extended_base_sample_shape(posterior=posterior)),
        )

    def _construct_base_samples(self, posterior) -> None:  # pragma: no cover - abstract
        raise NotImplementedError

    def _update_base_samples(self, posterior, base_sampler: "NormalMCSampler") -> None:
        """Redraw for ``posterior`` and overwrite the leading columns with ``base_sampler``'s draws.

        This is what keeps the baseline points' base samples fixed when q new points are appended
        (cached-Cholesky acquisition functions).  Single-output posteriors only: the points are the
        last dimension of the base samples.
        """
        self._instance_check(base_sampler=base_sampler)
        self._construct_base_samples(posterior=posterior)
        prev = base_sampler.base_samples
        if prev is None:
            return
        prev = prev.detach().clone()
        target = self._get_collapsed_shape(posterior=posterior)
        n_keep = prev.shape[-1]
        ns = len(self.sample_shape)
        lead = self.sample_shape + torch.Size([1] * (len(target) - prev.dim())) + prev.shape[ns:]
        self.base_samples[..., :n_keep] = prev.reshape(lead).expand(target[:-1] + prev.shape[-1:])


class IIDNormalSampler(NormalMCSampler):
    """iid N(0,1) base samples (``torch.randn`` under a temporary seed)."""

    def _construct_base_samples(self, posterior) -> None:
        target = self._get_collapsed_shape(posterior=posterior)
        if self.base_samples is None or self.base_samples.shape != target:
            with manual_seed(seed=self.seed):
                z = torch.randn(target, device=posterior.device, dtype=posterior.dtype)
            self.register_buffer("base_samples", z)
        if self.base_samples.device != posterior.device or self.base_samples.dtype != posterior.dtype:
            self.to(device=posterior.device, dtype=posterior.dtype)


class SobolQMCNormalSampler(NormalMCSampler):
    """Scrambled-Sobol N(0,1) base samples through the inverse CDF."""

    def _construct_base_samples(self, posterior, skip_leading: int = 0) -> None:
        """``skip_leading`` > 0: the caller overwrites the first ``skip_leading`` entries of the last dimension, so only
        the other Sobol dimensions are computed (``sobol_draw(columns=...)``); the skipped entries are left at zero."""
        target = self._get_collapsed_shape(posterior=posterior)
        if self.base_samples is None or self.base_samples.shape != target:
            dim = target[len(self.sample_shape) :].numel()
            if dim > SobolEngine.MAXDIM:
                raise UnsupportedError(
                    f"SobolQMCSampler only supports dimensions `q * o <= {SobolEngine.MAXDIM}`. Requested: {dim}"
                )
            n = self.sample_shape.numel()
            if skip_leading > 0:
                cols = torch.nonzero(torch.arange(dim) % target[-1] >= skip_leading).squeeze(-1)
                z = torch.zeros(n, dim, dtype=posterior.dtype)
                z[:, cols] = sobol_draw(dim, n, seed=self.seed, dtype=posterior.dtype, columns=cols,
                                        post=lambda u, sel: _normal_from_uniform(u))
                z = z.to(device=posterior.device)
            else:
                z = draw_sobol_normal_samples(d=dim, n=n, device=posterior.device, dtype=posterior.dtype, seed=self.seed)
            self.register_buffer("base_samples", z.view(target))
        self.to(device=posterior.device, dtype=posterior.dtype)

    def _update_base_samples(self, posterior, base_sampler: "NormalMCSampler") -> None:
        self._instance_check(base_sampler=base_sampler)
        prev = base_sampler.base_samples
        if prev is not None:
            self._construct_base_samples(posterior=posterior, skip_leading=prev.shape[-1])
        super()._update_base_samples(posterior=posterior, base_sampler=base_sampler)


def get_sampler(posterior, sample_shape: torch.Size, seed: Optional[int] = None) -> MCSampler:
    """Sobol when the collapsed base-sample dimension fits ``SobolEngine.MAXDIM``, iid otherwise."""
    lo, hi = posterior.batch_range
    bss = posterior.base_sample_shape
    dim = (bss[:lo] + bss[hi:]).numel()
    cls = SobolQMCNormalSampler if dim <= SobolEngine.MAXDIM else IIDNormalSampler
    return cls(sample_shape=sample_shape, seed=seed)
