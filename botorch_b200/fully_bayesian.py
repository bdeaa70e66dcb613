"""Fully Bayesian (SAAS) fit on the device potential.

What ``fit_fully_bayesian_model_nuts`` (botorch/fit.py:335-396) does is run pyro's NUTS on the potential energy of
``SaasPyroModel.sample`` (models/fully_bayesian.py:167-246) and load the thinned draws into the model.  The expensive part of
every leapfrog step is that potential: the log density of the training targets under ``outputscale * matern52 + noise I`` and
its gradient w.r.t. every hyperparameter.  Here that is ``bo_mll_value_grad`` (one blocked Cholesky + the ``W .* g`` GEMMs,
include/botorch_b200.h), evaluated for C chains as ONE batched call, and the sampler around it is a host-side NUTS that keeps
all state on the device (no host synchronisation except the tree-building decisions).

  * ``SaasPotential``                      -- pyro's ``potential_fn`` for the SAAS model, batched over chains
  * ``nuts_sample``                        -- multinomial NUTS with dual-averaging step size and diagonal mass adaptation
  * ``fit_fully_bayesian_model_nuts``      -- the reference's entry point and argument list

pyro is not in this image, so the SAMPLER cannot be pinned against the reference's draws (different RNG consumption even
with the same algorithm); the POTENTIAL is pinned at 1e-9 against oracle/saas.py (torch.distributions + autograd).

Flat layout of the unconstrained vector (length d + 4):  [ kernel_tausq, _kernel_inv_length_sq (d), mean, noise, outputscale ]
with the exp transform on every positive site (``biject_to(support)``) and its log-Jacobian in the energy.
"""
from __future__ import annotations

import ctypes
import math
from typing import Callable, Dict, Optional, Tuple

import torch
from torch import Tensor

from . import _lib, engine
from ._lib import check, ptr
from .models import MIN_INFERRED_NOISE_LEVEL, SaasFullyBayesianSingleTaskGP

_LOG_2_OVER_PI = math.log(2.0 / math.pi)


# Host mirrors of the prior terms inside saas_finalize_kernel (csrc/api.cu), pinned against torch.distributions by the CPU tests.
def _gamma_lp(x: Tensor, log_x: Tensor, a: float, b: float) -> Tensor:
    return a * math.log(b) + (a - 1.0) * log_x - b * x - math.lgamma(a)


def _half_cauchy_lp(x: Tensor, scale: float) -> Tensor:
    return _LOG_2_OVER_PI - math.log(scale) - torch.log1p((x / scale) ** 2)


class SaasPotential:
    """Potential energy (negative log joint in unconstrained space) of the SAAS model for ``num_chains`` parameter vectors."""

    def __init__(self, train_X: Tensor, train_Y: Tensor, num_chains: int = 1, graph: bool = True) -> None:
        if not train_X.is_cuda:
            raise RuntimeError("SaasPotential lives on the GPU (no CPU path); move train_X / train_Y to a cuda device")
        self.X = train_X.detach().to(torch.float64).contiguous()
        self.y = train_Y.detach().to(torch.float64).reshape(-1).contiguous()
        self.d = self.X.shape[-1]
        self.dim = self.d + 4
        self.num_chains = int(num_chains)
        self._state = engine.MLLState(self.X, self.y, "matern52", with_outputscale=True, M=self.num_chains)
        self._wild = torch.zeros(self.num_chains, dtype=torch.int32, device=self.X.device)
        self.evaluations = 0
        # The whole evaluation (bo_saas_potential: transforms, bo_mll_value_grad, prior terms, chain rule) is a fixed launch sequence without
        # host synchronisation: after two eager calls it is captured once and replayed as one CUDA graph per leapfrog step.
        self._use_graph, self._graph, self._eager_calls = bool(graph), None, 0

    def constrain(self, u: Tensor) -> Dict[str, Tensor]:
        """Unconstrained [C x (d+4)] -> the dict ``postprocess_mcmc_samples`` produces (fully_bayesian.py:248-265)."""
        d = self.d
        tausq, ils = u[:, 0].exp(), u[:, 1 : 1 + d].exp()
        # "noise" is the raw Gamma site, as pyro records it: the likelihood inside NUTS uses 1e-4 + noise
        # (fully_bayesian.py:215-221) but the stored draw is the site value, and `load_mcmc_samples` clamps it from below
        # (`clamp_min(MIN_INFERRED_NOISE_LEVEL)`, fully_bayesian.py:298) -- max(g, 1e-4), not g + 1e-4.  Same here.
        return {"mean": u[:, 1 + d], "noise": u[:, 2 + d].exp(), "outputscale": u[:, 3 + d].exp(),
                "lengthscale": (tausq.unsqueeze(-1) * ils).rsqrt()}  # fmt: skip

    def __call__(self, u: Tensor) -> Tuple[Tensor, Tensor]:
        """u: [C x (d+4)] -> (energy [C], d energy / d u [C x (d+4)]).  A chain whose covariance cannot be factored or whose
        energy is not finite gets energy +inf and a zero gradient -- a divergent transition for the sampler, which is how
        pyro treats the LinAlgError / PositiveDefinite exceptions the reference registers (fully_bayesian.py:67-78).  The
        other chains of the batch are unaffected."""
        if u.shape != (self.num_chains, self.dim):
            raise ValueError(f"expected u of shape {(self.num_chains, self.dim)}, got {tuple(u.shape)}")
        self.evaluations += 1
        u = u.detach().to(torch.float64)
        if not self._use_graph:
            return self._evaluate(u)
        if self._graph is None:
            if self._eager_calls < 2:  # warm-up: one-time kernel attributes and allocator pools settle outside the capture
                self._eager_calls += 1
                return self._evaluate(u)
            self._u = u.clone()
            torch.cuda.synchronize(u.device)
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._out = self._evaluate(self._u)
        self._u.copy_(u)
        self._graph.replay()
        return self._out[0].clone(), self._out[1].clone()

    @property
    def capture_safe_evaluate(self) -> PotentialFn:
        """The evaluation as a plain launch sequence (no replay of this object's own graph): what `nuts_sample` captures
        inside its per-depth subtree graphs."""
        return self._evaluate

    def _evaluate(self, u: Tensor) -> Tuple[Tensor, Tensor]:
        """One ``bo_saas_potential``: transform kernel -> bo_mll_value_grad -> closed-form chain-rule kernel.  No host
        synchronisation, no allocation inside the library: the launch sequence is graph-capturable as it is."""
        st, gp = self._state, self._state.gp
        energy = torch.empty(self.num_chains, dtype=torch.float64, device=u.device)
        grad_u = torch.empty(self.num_chains, self.dim, dtype=torch.float64, device=u.device)
        u = u.contiguous()
        gp.c.jitter = None  # one plain factorisation: a covariance that cannot be factored is a divergent transition
        check(_lib.lib().bo_saas_potential(gp.ctx, ctypes.byref(gp.c), ptr(u), st.work.data_ptr(), st.work.numel(), ptr(st.value),
                                           ptr(st.grad), ptr(self._wild), ptr(energy), ptr(grad_u), ptr(gp.info),
                                           _lib.stream_ptr(u.device)))  # fmt: skip
        return energy, grad_u


# ---- NUTS over a batch of chains ---------------------------------------------------------------------------------------------
# Multinomial NUTS (Hoffman & Gelman 2014 with Betancourt's multinomial sampling and the generalised U-turn criterion, the
# variant pyro and Stan run), written for C chains in LOCKSTEP: every chain doubles its tree at the same time, a chain whose
# tree has terminated rides along masked out, so each leapfrog step of all chains is one batched potential call.
PotentialFn = Callable[[Tensor], Tuple[Tensor, Tensor]]
_MAX_DELTA_ENERGY = 1000.0  # divergence threshold (pyro's NUTS._max_sliced_energy)


class _Tree:
    """Per-chain subtree summary.  `in` is the end the subtree was grown from (only its momentum is needed: growth continues
    from the far end `out`); momenta keep their absolute orientation (backward growth integrates with a negative step), so the U-turn test is symmetric in the two ends."""

    __slots__ = ("r_in", "z_out", "r_out", "g_out", "z_prop", "u_prop", "g_prop", "log_w", "rho", "stop",
                 "diverged", "sum_accept", "leaves")  # fmt: skip

    def select(self, keep_self: Tensor, other: "_Tree") -> "_Tree":
        out = _Tree()
        for name in self.__slots__:
            a, b = getattr(self, name), getattr(other, name)
            m = keep_self if a.dim() == 1 else keep_self.unsqueeze(-1)
            setattr(out, name, torch.where(m, a, b))
        return out


def _kinetic(r: Tensor, minv: Tensor) -> Tensor:
    return 0.5 * (r * r * minv).sum(-1)


def _leaf(potential: PotentialFn, z, r, g, step, minv, h0) -> _Tree:
    s = step.unsqueeze(-1)
    r_half = r - 0.5 * s * g
    z_new = z + s * minv * r_half
    u_new, g_new = potential(z_new)
    r_new = r_half - 0.5 * s * g_new
    delta = u_new + _kinetic(r_new, minv) - h0
    delta = torch.where(torch.isfinite(delta), delta, torch.full_like(delta, float("inf")))
    t = _Tree()
    t.z_out = t.z_prop = z_new
    t.r_in = t.r_out = r_new
    t.g_out = t.g_prop = g_new
    t.u_prop = u_new
    t.log_w = -delta
    t.rho = r_new
    t.diverged = delta > _MAX_DELTA_ENERGY
    t.stop = t.diverged.clone()
    t.sum_accept = torch.exp(torch.clamp(-delta, max=0.0))
    t.leaves = torch.ones_like(delta)
    return t


def _turning(rho: Tensor, r_a: Tensor, r_b: Tensor, minv: Tensor) -> Tensor:
    return ((rho * minv * r_a).sum(-1) <= 0) | ((rho * minv * r_b).sum(-1) <= 0)


def _build(potential: PotentialFn, depth: int, z, r, g, step, minv, h0, R: Tensor, base: int = 0) -> _Tree:
    """Subtree of 2**depth leapfrog steps.  R [2**depth x C]: uniform draws, one row per merge (2**depth - 1 of them: row
    `base` for this node, the next 2**(depth-1) - 1 for the first half, the rest for the second)."""
    if depth == 0:
        return _leaf(potential, z, r, g, step, minv, h0)
    half = 2 ** (depth - 1)
    first = _build(potential, depth - 1, z, r, g, step, minv, h0, R, base + 1)
    second = _build(potential, depth - 1, first.z_out, first.r_out, first.g_out, step, minv, h0, R, base + half)
    both = _Tree()
    both.r_in = first.r_in
    both.z_out, both.r_out, both.g_out = second.z_out, second.r_out, second.g_out
    both.log_w = torch.logaddexp(first.log_w, second.log_w)
    take2 = R[base] < torch.exp(second.log_w - both.log_w)
    both.z_prop = torch.where(take2.unsqueeze(-1), second.z_prop, first.z_prop)
    both.g_prop = torch.where(take2.unsqueeze(-1), second.g_prop, first.g_prop)
    both.u_prop = torch.where(take2, second.u_prop, first.u_prop)
    both.rho = first.rho + second.rho
    both.diverged = second.diverged
    both.stop = second.stop | _turning(both.rho, both.r_in, both.r_out, minv)
    both.sum_accept = first.sum_accept + second.sum_accept
    both.leaves = first.leaves + second.leaves
    return first.select(first.stop, both)  # a subtree that already stopped ignores everything grown after it


class _GraphedTrees:
    """`_build` for each depth captured once into a CUDA graph over static buffers (CUDA tensors and a potential that offers
    a capture-safe evaluation): growing a subtree of 2**depth leapfrog steps is then ONE graph launch instead of
    ~100 * 2**depth kernel launches issued from Python."""

    def __init__(self, evaluate: PotentialFn, like: Tensor) -> None:
        self.evaluate, self.graphs = evaluate, {}
        self.z, self.r, self.g, self.minv = (torch.zeros_like(like) for _ in range(4))
        self.step, self.h0 = torch.zeros_like(like[:, 0]), torch.zeros_like(like[:, 0])

    def build(self, depth: int, z, r, g, step, minv, h0, R: Tensor) -> _Tree:
        for dst, src in ((self.z, z), (self.r, r), (self.g, g), (self.step, step), (self.minv, minv), (self.h0, h0)):
            dst.copy_(src)
        if depth not in self.graphs:
            R_static = R.clone()
            torch.cuda.synchronize(z.device)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                tree = _build(self.evaluate, depth, self.z, self.r, self.g, self.step, self.minv, self.h0, R_static)
            self.graphs[depth] = (graph, R_static, tree)
        graph, R_static, tree = self.graphs[depth]
        R_static.copy_(R)
        graph.replay()
        return tree  # fields alias the graph's output buffers: consumed (torch.where into new tensors) before the next replay


class _DualAveraging:
    """Nesterov dual averaging of log step size per chain (Hoffman & Gelman 2014, section 3.2; pyro's constants)."""

    def __init__(self, eps0: Tensor, target: float = 0.8, gamma: float = 0.05, t0: float = 10.0, kappa: float = 0.75):
        self.target, self.gamma, self.t0, self.kappa = target, gamma, t0, kappa
        self.restart(eps0)

    def restart(self, eps0: Tensor) -> None:
        self.mu = torch.log(10.0 * eps0)
        self.h_bar = torch.zeros_like(eps0)
        self.log_eps_bar = torch.zeros_like(eps0)
        self.t = 0

    def update(self, accept: Tensor) -> Tensor:
        self.t += 1
        w = 1.0 / (self.t + self.t0)
        self.h_bar = (1.0 - w) * self.h_bar + w * (self.target - accept)
        log_eps = self.mu - math.sqrt(self.t) / self.gamma * self.h_bar
        eta = self.t ** (-self.kappa)
        self.log_eps_bar = eta * log_eps + (1.0 - eta) * self.log_eps_bar
        return torch.exp(log_eps)

    def final(self) -> Tensor:
        return torch.exp(self.log_eps_bar)


def _reasonable_step(potential: PotentialFn, z, u, g, minv, gen, eps0: float = 0.1, tries: int = 24) -> Tensor:
    """Per chain: double / halve the step until the one-step acceptance probability crosses 1/2 (H&G Algorithm 4)."""
    C = z.shape[0]
    eps = torch.full((C,), eps0, dtype=z.dtype, device=z.device)
    r = torch.randn(z.shape, generator=gen, device=z.device, dtype=z.dtype) / minv.sqrt()
    h0 = u + _kinetic(r, minv)
    direction = None
    done = torch.zeros(C, dtype=torch.bool, device=z.device)
    for _ in range(tries):
        leaf = _leaf(potential, z, r, g, eps, minv, h0)
        up = leaf.log_w > math.log(0.5)  # acceptance above 1/2: the step can grow
        if direction is None:
            direction = up
        done = done | (up != direction)
        if bool(done.all()):
            break
        eps = torch.where(done, eps, torch.where(direction, eps * 2.0, eps * 0.5))
    return eps


class _Welford:
    def __init__(self, like: Tensor) -> None:
        self.n, self.mean, self.m2 = 0, torch.zeros_like(like), torch.zeros_like(like)

    def add(self, x: Tensor) -> None:
        self.n += 1
        delta = x - self.mean
        self.mean = self.mean + delta / self.n
        self.m2 = self.m2 + delta * (x - self.mean)

    def variance(self) -> Tensor:  # Stan's regularisation towards the unit metric
        var = self.m2 / max(self.n - 1, 1)
        return (self.n / (self.n + 5.0)) * var + 1e-3 * (5.0 / (self.n + 5.0))


def _adaptation_windows(warmup: int) -> Tuple[int, list]:
    """(start of the metric phase, ends of its doubling windows); the last 10% adapt the step size only (Stan's schedule)."""
    if warmup < 40:
        return warmup, []
    start, end = int(0.15 * warmup), int(0.9 * warmup)
    ends, size, pos = [], 25, start
    while pos + 2 * size + size <= end:
        pos += size
        ends.append(pos)
        size *= 2
    ends.append(end)
    return start, ends


def nuts_sample(potential: PotentialFn, init: Tensor, num_samples: int, warmup_steps: int, max_tree_depth: int = 6,
                target_accept: float = 0.8, seed: Optional[int] = None, adapt_mass: bool = True,
                graph: Optional[bool] = None):  # fmt: skip
    """Run C chains of NUTS in lockstep.  init: [C x D] -> (samples [num_samples x C x D], stats).  ``potential`` maps
    [C x D] to (energy [C], gradient [C x D]); +inf energy marks a divergent point.  With ``graph`` (default: whenever the
    potential offers ``capture_safe_evaluate`` and lives on a CUDA device) each subtree is one CUDA-graph launch; the draws
    are the same with and without graphs."""
    z = init.detach().clone()
    dev, dt = z.device, z.dtype
    C = z.shape[0]
    gen = torch.Generator(device=dev)
    gen.manual_seed(torch.seed() % (2**31) if seed is None else int(seed))
    u, g = potential(z)
    if not bool(torch.isfinite(u).all()):
        raise RuntimeError("NUTS needs a finite potential energy at every initial point")
    capture_safe = getattr(potential, "capture_safe_evaluate", None)
    if graph is None:
        graph = capture_safe is not None and z.is_cuda
    if graph and (capture_safe is None or not z.is_cuda):
        raise ValueError("graph=True needs CUDA tensors and a potential with `capture_safe_evaluate`")
    trees = _GraphedTrees(capture_safe, z) if graph else None
    minv = torch.ones_like(z)
    eps = _reasonable_step(potential, z, u, g, minv, gen)
    averaging = _DualAveraging(eps, target=target_accept)
    metric_start, window_ends = _adaptation_windows(warmup_steps) if adapt_mass else (warmup_steps, [])
    welford = _Welford(z)
    samples = torch.empty(num_samples, C, z.shape[1], dtype=dt, device=dev)
    stats = {"accept_prob": [], "tree_depth": [], "divergent": torch.zeros(C, dtype=torch.long, device=dev), "leapfrogs": 0}
    for it in range(warmup_steps + num_samples):
        r0 = torch.randn(z.shape, generator=gen, device=dev, dtype=dt) / minv.sqrt()
        h0 = u + _kinetic(r0, minv)
        z_minus = z_plus = z
        r_minus = r_plus = r0
        g_minus = g_plus = g
        rho, log_w = r0.clone(), torch.zeros_like(u)
        active = torch.ones(C, dtype=torch.bool, device=dev)
        sum_accept, leaves = torch.zeros_like(u), torch.zeros_like(u)
        depth_reached = torch.zeros(C, dtype=torch.long, device=dev)
        for depth in range(max_tree_depth):
            fwd = torch.rand(C, generator=gen, device=dev, dtype=dt) < 0.5
            f2 = fwd.unsqueeze(-1)
            R = torch.rand(2**depth, C, generator=gen, device=dev, dtype=dt)
            edge = (torch.where(f2, z_plus, z_minus), torch.where(f2, r_plus, r_minus), torch.where(f2, g_plus, g_minus),
                    torch.where(fwd, eps, -eps), minv, h0, R)  # fmt: skip
            if trees is not None:
                tree = trees.build(depth, *edge)
                if hasattr(potential, "evaluations"):
                    potential.evaluations += 2**depth
            else:
                tree = _build(potential, depth, *edge)
            stats["leapfrogs"] += 2**depth
            sum_accept = sum_accept + torch.where(active, tree.sum_accept, torch.zeros_like(u))
            leaves = leaves + torch.where(active, tree.leaves, torch.zeros_like(u))
            depth_reached = depth_reached + active.long()
            if it >= warmup_steps:
                stats["divergent"] += (active & tree.diverged).long()
            valid = active & ~tree.stop
            # biased progressive sampling between the old tree and the new subtree (Betancourt 2017, A.3.2)
            take = valid & (torch.rand(C, generator=gen, device=dev, dtype=dt) < torch.exp(tree.log_w - log_w))
            t2 = take.unsqueeze(-1)
            z, g, u = torch.where(t2, tree.z_prop, z), torch.where(t2, tree.g_prop, g), torch.where(take, tree.u_prop, u)
            grow_f, grow_b = (valid & fwd).unsqueeze(-1), (valid & ~fwd).unsqueeze(-1)
            z_plus, r_plus, g_plus = (torch.where(grow_f, tree.z_out, z_plus), torch.where(grow_f, tree.r_out, r_plus),
                                      torch.where(grow_f, tree.g_out, g_plus))  # fmt: skip
            z_minus, r_minus, g_minus = (torch.where(grow_b, tree.z_out, z_minus), torch.where(grow_b, tree.r_out, r_minus),
                                         torch.where(grow_b, tree.g_out, g_minus))  # fmt: skip
            log_w = torch.where(valid, torch.logaddexp(log_w, tree.log_w), log_w)
            rho = torch.where(valid.unsqueeze(-1), rho + tree.rho, rho)
            active = valid & ~_turning(rho, r_minus, r_plus, minv)
            if not bool(active.any()):
                break
        accept = sum_accept / leaves.clamp_min(1.0)
        if it < warmup_steps:
            eps = averaging.update(accept)
            if window_ends and it >= metric_start:
                welford.add(z)
                if it + 1 == window_ends[0]:
                    window_ends.pop(0)
                    minv = welford.variance()
                    welford = _Welford(z)
                    eps = _reasonable_step(potential, z, u, g, minv, gen, eps0=1.0)
                    averaging.restart(eps)
            if it + 1 == warmup_steps:
                eps = averaging.final()
        else:
            samples[it - warmup_steps] = z
            stats["accept_prob"].append(accept)
            stats["tree_depth"].append(depth_reached)
    stats["step_size"], stats["inverse_mass"] = eps, minv
    if stats["accept_prob"]:
        stats["accept_prob"] = torch.stack(stats["accept_prob"]).mean(0)
        stats["tree_depth"] = torch.stack(stats["tree_depth"]).double().mean(0)
    return samples, stats


def fit_fully_bayesian_model_nuts(model: SaasFullyBayesianSingleTaskGP, max_tree_depth: int = 6, warmup_steps: int = 512,
                                  num_samples: int = 256, thinning: int = 16, disable_progbar: bool = False,
                                  jit_compile: bool = False, num_chains: int = 1, seed: Optional[int] = None) -> None:  # fmt: skip
    """botorch/fit.py:335-396: NUTS on the SAAS potential, every ``thinning``-th draw loaded into ``model``.

    ``num_chains`` (not in the reference, which runs one chain) runs that many chains as ONE batched potential; each
    contributes ``num_samples // num_chains`` draws, so the retained count stays ``num_samples / thinning``.  Initial points
    are uniform in (-2, 2) in the unconstrained space (pyro's ``init_to_uniform`` default).  ``disable_progbar`` and
    ``jit_compile`` are accepted for signature compatibility and have no effect."""
    if not isinstance(model, SaasFullyBayesianSingleTaskGP):
        raise TypeError("fit_fully_bayesian_model_nuts fits a SaasFullyBayesianSingleTaskGP")
    model.train()
    X = model.train_inputs[0]
    potential = SaasPotential(X, model.train_targets, num_chains=num_chains)
    gen = torch.Generator(device=X.device)
    gen.manual_seed(torch.seed() % (2**31) if seed is None else int(seed))
    init = None
    for _ in range(20):  # redraw the starting points whose energy is not finite
        trial = 4.0 * torch.rand(num_chains, potential.dim, generator=gen, device=X.device, dtype=torch.float64) - 2.0
        init = trial if init is None else torch.where(torch.isfinite(potential(init)[0]).unsqueeze(-1), init, trial)
        if bool(torch.isfinite(potential(init)[0]).all()):
            break
    per_chain = max(num_samples // num_chains, 1)
    samples, stats = nuts_sample(potential, init, per_chain, warmup_steps, max_tree_depth=max_tree_depth,
                                 seed=None if seed is None else seed + 1)  # fmt: skip
    kept = samples.transpose(0, 1).reshape(-1, potential.dim)[::thinning]  # chain-major, then thinned like the reference
    model.load_mcmc_samples(potential.constrain(kept))
    model.eval()
    model.nuts_stats = {**stats, "potential_evaluations": potential.evaluations}
