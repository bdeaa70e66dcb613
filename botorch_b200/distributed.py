"""Sharding the multi-start batch across the GPUs of one box (one process per GPU, torch.distributed).

The reference has no multi-device code; its seam is the ``batch_limit`` / ``init_batch_limit``
chunking of ``optimize_acqf`` (botorch/optim/optimize.py:268-317, botorch/optim/initializers.py:411-423): chunks of
t-batches never interact.  Here each chunk goes to one rank:

  * training-set factors and the qLogNEI baseline state are computed ONCE on rank 0 and broadcast (NCCL over
    NVLink) -- the only bulk transfer;
  * raw samples: rank r evaluates the contiguous block ``shard_range(R, r, G)``; the R values are all-gathered
    (R doubles) so every rank performs the identical host-side selection (global indices preserved => the
    argmax of optim/initializers.py:944 is bit-identical to the single-GPU run);
  * restarts: rank r optimises its block (== the reference with ``batch_limit = B/G``); per-rank
    (best value, global index, candidate) are all-gathered and the winner is the largest value with the lowest
    index on ties, i.e. ``torch.argmax`` first-occurrence semantics (optim/optimize.py:385).

There is no per-iteration collective.  All functions degrade to the single-process case when torch.distributed is
not initialised, and work on the gloo backend for the CPU tests of the host logic.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist
from torch import Tensor


_GROUP = None  # process group the sharding functions run over (None: the default group)


class use_group:
    """Context manager: run the sharding functions of this module over a sub-group of the job's ranks (e.g. the first N
    GPUs of the box).  Only the members may call them inside the context; ``src`` arguments are ranks WITHIN the group."""

    def __init__(self, group) -> None:
        self.group, self.prev = group, None

    def __enter__(self):
        global _GROUP
        self.prev, _GROUP = _GROUP, self.group
        return self.group

    def __exit__(self, *exc) -> None:
        global _GROUP
        _GROUP = self.prev


def world() -> Tuple[int, int]:
    """(rank, world_size) in the active group; (0, 1) without an initialised process group."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(_GROUP), dist.get_world_size(_GROUP)
    return 0, 1


def _backend() -> str:
    return dist.get_backend(_GROUP)


def _bcast(t: Tensor, src: int) -> None:
    dist.broadcast(t, src=src if _GROUP is None else dist.get_global_rank(_GROUP, src), group=_GROUP)


def _all_gather(out: List[Tensor], t: Tensor) -> None:
    dist.all_gather(out, t, group=_GROUP)


def shard_range(total: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of ``total`` units for ``rank``; the first ``total % world_size`` ranks get one more."""
    base, extra = divmod(total, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def warm_up(device=None) -> None:
    """Bring the communicator up (NCCL creates its channels lazily on the first collective: ~1 s on an 8-GPU box) so that
    the first acquisition does not pay for it.  Call once after ``init_process_group``."""
    _, ws = world()
    if ws == 1:
        return
    t = torch.zeros(1, dtype=torch.float64, device=device if _backend() == "nccl" else "cpu")
    dist.all_reduce(t, group=_GROUP)
    _bcast(t, 0)
    if t.is_cuda:
        torch.cuda.synchronize(t.device)


def shared_seed(seed: Optional[int], device=None) -> Optional[int]:
    """The seed every rank must agree on.  A given seed is returned as is; ``None`` means "draw one", which has to happen on
    ONE rank (rank 0, from its global RNG like the reference's samplers do) and be broadcast -- otherwise every rank would
    scramble its own Sobol sequence and the gathered values would belong to different points."""
    rank, ws = world()
    if seed is not None or ws == 1:
        return seed
    t = torch.randint(0, 2**31 - 1, (1,), dtype=torch.int64)
    if _backend() == "nccl":
        t = t.to(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
    _bcast(t, 0)
    return int(t.item())


def broadcast_buffers(buffers: List[Tensor], src: int = 0) -> int:
    """Broadcast every buffer in place from ``src``; returns the payload in bytes.

    The buffers are packed into ONE contiguous allocation per dtype and sent with a single collective (one NCCL launch
    instead of ~25 blocking ones; the two extra device copies cost 2 x payload / HBM bandwidth, ~0.15 ms for the 0.4 GB of
    config C3)."""
    nbytes = sum(b.numel() * b.element_size() for b in buffers)
    rank, ws = world()
    if ws == 1 or not buffers:
        return nbytes
    by_dtype = {}
    for b in buffers:
        by_dtype.setdefault(b.dtype, []).append(b)
    for group in by_dtype.values():
        if len(group) == 1:
            _bcast(group[0], src)
            continue
        total = sum(b.numel() for b in group)
        flat = torch.empty(total, dtype=group[0].dtype, device=group[0].device)
        if rank == src:
            torch.cat([b.reshape(-1) for b in group], out=flat)
        _bcast(flat, src)
        if rank != src:
            off = 0
            for b in group:
                b.copy_(flat[off : off + b.numel()].view_as(b))
                off += b.numel()
    return nbytes


def replicate_model(model, src: int = 0, factor_mode: str = "broadcast") -> int:
    """Factor the training set on ``src`` only and broadcast the factors (K2 of SURVEY.md: once per model).
    ``factor_mode="recompute"``: every rank factors its own copy instead (the kernels are deterministic, so the factors
    are bit-identical on every rank; no transfer at all -- cheaper than any broadcast while n is a few thousand)."""
    rank, ws = world()
    if ws == 1 or factor_mode == "recompute":
        model.gp_state
        return 0
    model.eval()
    model._state = model._build_state(setup=(rank == src))
    return broadcast_buffers(model._state.factor_buffers(), src)


def make_qlognei(model, X_baseline: Tensor, sampler=None, src: int = 0, factor_mode: str = "broadcast", **kwargs):
    """qLogNoisyExpectedImprovement whose baseline state is computed on ``src`` and broadcast.

    Returns (acquisition function, broadcast bytes).  Base samples are drawn on every rank from the same seed
    (host SobolEngine), so only factors travel; with ``sampler=None`` the seed is drawn on rank 0 and broadcast.
    """
    from .acquisition import DEFAULT_MC_SAMPLES, qLogNoisyExpectedImprovement
    from .sampling import SobolQMCNormalSampler

    rank, ws = world()
    nbytes = replicate_model(model, src, factor_mode)
    if ws > 1 and sampler is None:
        sampler = SobolQMCNormalSampler(sample_shape=torch.Size([DEFAULT_MC_SAMPLES]),
                                        seed=shared_seed(None, X_baseline.device))  # fmt: skip
    elif ws > 1 and getattr(sampler, "seed", None) is not None:
        # a sampler constructed without a seed drew its own on every rank: adopt rank 0's
        t = torch.tensor([int(sampler.seed)], dtype=torch.int64)
        if _backend() == "nccl":
            t = t.to(X_baseline.device)
        _bcast(t, src)
        sampler.seed = int(t.item())
    local_setup = ws == 1 or rank == src or factor_mode == "recompute"
    acqf = qLogNoisyExpectedImprovement(model, X_baseline, sampler=sampler, _factor_setup=local_setup, **kwargs)
    if ws > 1 and factor_mode != "recompute":
        nbytes += broadcast_buffers(acqf._nei.factor_buffers(), src)
        acqf._refresh_baseline_views()
    return acqf, nbytes


def all_gather_blocks(local: Tensor, total: int) -> Tensor:
    """Concatenate per-rank contiguous blocks (sizes given by ``shard_range``) into the full vector on every rank."""
    rank, ws = world()
    if ws == 1:
        return local
    sizes = [shard_range(total, r, ws) for r in range(ws)]
    width = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros(width, *local.shape[1:], dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = [torch.empty_like(pad) for _ in range(ws)]
    _all_gather(out, pad)
    return torch.cat([o[: hi - lo] for o, (lo, hi) in zip(out, sizes)], dim=0)


def sharded_raw_sample_values(acq_function, X_rnd: Tensor, device, batch_limit: Optional[int] = None,
                              total: Optional[int] = None) -> Tensor:  # fmt: skip
    """Acquisition values of all raw q-batches; each rank evaluates its contiguous block (global order kept).  ``X_rnd`` is
    the whole set (every rank holds it and slices its block) or, with ``total`` given, this rank's block of ``total``."""
    from .optim import evaluate_raw_samples

    rank, ws = world()
    if total is None:
        total = X_rnd.shape[0]
        lo, hi = shard_range(total, rank, ws)
        X_rnd = X_rnd[lo:hi]
    y_local = evaluate_raw_samples(acq_function, X_rnd, device, batch_limit)
    if ws == 1:
        return y_local
    backend = _backend()
    y_dev = y_local.to(device) if backend == "nccl" else y_local
    return all_gather_blocks(y_dev, total).cpu()


def gather_best(values: Tensor, candidates: Tensor, index_offset: int) -> Tuple[Tensor, Tensor, int]:
    """Global best over all ranks' (values [b], candidates [b x q x d]).

    Each rank contributes its local argmax (first occurrence), the winner is the largest value, lowest global index
    on ties.  Returns (candidate q x d, value, global index) on every rank.
    """
    rank, ws = world()
    if values.numel() == 0:  # a rank whose shard is empty (more ranks than restarts) never wins
        best_val = torch.full((1,), -float("inf"), dtype=torch.float64, device=values.device)
        best_idx = torch.full((1,), torch.iinfo(torch.int64).max, dtype=torch.int64, device=values.device)
        best_cand = torch.zeros(candidates.shape[1:], dtype=torch.float64, device=values.device)
        li = -1
    else:
        li = int(torch.argmax(values.view(-1)))
        best_val = values.view(-1)[li].reshape(1).to(torch.float64)
        best_idx = torch.tensor([index_offset + li], dtype=torch.int64, device=values.device)
        best_cand = candidates[li].to(torch.float64).contiguous()
    if ws == 1:
        return candidates[li], values.view(-1)[li], int(best_idx)
    vals = [torch.empty_like(best_val) for _ in range(ws)]
    idxs = [torch.empty_like(best_idx) for _ in range(ws)]
    cands = [torch.empty_like(best_cand) for _ in range(ws)]
    _all_gather(vals, best_val)
    _all_gather(idxs, best_idx)
    _all_gather(cands, best_cand)
    vals_t, idxs_t = torch.cat(vals), torch.cat(idxs)
    top = vals_t.max()
    tied = torch.nonzero(vals_t == top).view(-1)
    win = int(tied[torch.argmin(idxs_t[tied])])
    return cands[win].to(candidates.dtype), vals_t[win].to(values.dtype), int(idxs_t[win])


TS_NOISE_BLOCK = 16384  # candidates per base-sample block: the noise of candidate i depends on (seed, draw, i // block) only


def ts_base_samples(lo: int, hi: int, draw: int, seed: int, device) -> Tensor:
    """Standard-normal base samples of the candidates [lo, hi) for one Thompson draw, identical for every sharding:
    block k of ``TS_NOISE_BLOCK`` candidates is generated from its own counter-derived seed."""
    out = torch.empty(hi - lo, dtype=torch.float64, device=device)
    gen = torch.Generator(device=device)
    k0, k1 = lo // TS_NOISE_BLOCK, (hi - 1) // TS_NOISE_BLOCK
    for k in range(k0, k1 + 1):
        gen.manual_seed((int(seed) * 1_000_003 + draw) * 1_000_003 + k)
        z = torch.randn(TS_NOISE_BLOCK, dtype=torch.float64, device=device, generator=gen)
        a, b = max(lo, k * TS_NOISE_BLOCK), min(hi, (k + 1) * TS_NOISE_BLOCK)
        out[a - lo : b - lo] = z[a - k * TS_NOISE_BLOCK : b - k * TS_NOISE_BLOCK]
    return out


class _ShardScorer:
    """Marginal posterior of the shard [lo, hi) of a replicated candidate set, and per-draw Thompson picks inside it."""

    def __init__(self, sampler, X: Tensor, lo: int, hi: int) -> None:
        from . import _lib

        st = sampler.model.gp_state
        if st.M != 1:
            raise NotImplementedError("sharded Thompson sampling is implemented for single (non-ensemble) models")
        self.st, self.lo, self.hi, self.dev = st, lo, hi, st.device
        self.lib = _lib
        if hi > lo:
            mean, var = sampler.marginal_posterior(X[lo:hi].to(torch.float64).contiguous())
            self.mean, self.var = mean[0].clone().contiguous(), var[0].contiguous()
        self.val = torch.empty(1, dtype=torch.float64, device=self.dev)
        self.idx = torch.empty(1, dtype=torch.int64, device=self.dev)

    def pick(self, draw: int, seed: int):
        """(value [1], global index [1]) of the best candidate of this shard for one draw; (-inf, int64 max) if empty."""
        if self.hi <= self.lo:
            return (torch.full((1,), -float("inf"), dtype=torch.float64, device=self.dev),
                    torch.full((1,), torch.iinfo(torch.int64).max, dtype=torch.int64, device=self.dev))  # fmt: skip
        z = ts_base_samples(self.lo, self.hi, draw, seed, self.dev)
        lib = self.lib
        lib.check(lib.lib().bo_ts_argmax(self.st.ctx, 1, self.hi - self.lo, self.mean.data_ptr(), self.var.data_ptr(),
                                         z.data_ptr(), self.val.data_ptr(), self.idx.data_ptr(),
                                         lib.stream_ptr(self.dev)))  # fmt: skip
        return self.val.clone(), self.idx + self.lo

    def exclude(self, global_idx: int) -> None:
        if self.lo <= global_idx < self.hi:
            self.mean[global_idx - self.lo] = -float("inf")


def _winner(vals: Tensor, idxs: Tensor) -> Tensor:
    """Across ranks (dim 0): index of the largest value, lowest global index on ties (torch.argmax semantics)."""
    top = vals.max(dim=0).values
    masked = torch.where(vals == top, idxs, torch.full_like(idxs, torch.iinfo(torch.int64).max))
    return masked.min(dim=0).values


def pathwise_shard_scores(sampler, X: Tensor, num_samples: int, seed: int, lo: int, hi: int):
    """(best value, global index) per sample path over the candidate block [lo, hi).  Every rank draws the SAME paths: the
    host generators run from ``seed`` (state restored afterwards), so the S functions are identical everywhere and only
    the candidates are sharded."""
    from .generation import MatheronPath

    state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    try:
        paths = MatheronPath(sampler.model, torch.Size([num_samples]), sampler.num_features, sampler.chunk_size)
    finally:
        torch.random.set_rng_state(state)
    vals = paths(X[lo:hi])
    v, i = vals.max(dim=-1)  # first occurrence on ties
    return v, i + lo


def thompson_sample_sharded(sampler, X: Tensor, num_samples: int = 1, seed: int = 0, shard=None) -> Tensor:
    """``MaxPosteriorSampling`` (marginal form, SURVEY.md row T1 / 8e "C5: shard N") over a candidate set sharded across
    the process group: rank r scores the contiguous block ``shard_range(N, r, G)``; (value, global index) pairs are
    all-gathered -- once per call with replacement, once per draw without -- and the winner is the largest value,
    lowest index on ties.  Base samples are keyed by global candidate position, so the picks do not depend on the
    number of ranks.  ``X`` (N x d) is replicated; ``shard=(rank, world_size)`` overrides the process group (tests).
    """
    rank, ws = world() if shard is None else shard
    gather = shard is None and ws > 1
    lo, hi = shard_range(X.shape[0], rank, ws)

    def all_ranks(v: Tensor, i: Tensor):
        if not gather:
            return v.unsqueeze(0), i.unsqueeze(0)
        gv = [torch.empty_like(v) for _ in range(ws)]
        gi = [torch.empty_like(i) for _ in range(ws)]
        _all_gather(gv, v)
        _all_gather(gi, i)
        return torch.stack(gv), torch.stack(gi)

    if getattr(sampler, "pathwise", False):  # joint draws: one function per sample, argmax over all shards
        if not sampler.replacement:
            from .exceptions import UnsupportedError

            raise UnsupportedError("sharded pathwise Thompson sampling draws with replacement")
        V, I = all_ranks(*pathwise_shard_scores(sampler, X, num_samples, seed, lo, hi))
        return X[_winner(V, I)]
    scorer = _ShardScorer(sampler, X, lo, hi)
    if sampler.replacement:
        picks = [scorer.pick(s, seed) for s in range(num_samples)]
        V, I = all_ranks(torch.cat([p[0] for p in picks]), torch.cat([p[1] for p in picks]))
        return X[_winner(V, I)]
    chosen = []
    for s in range(num_samples):  # sequential: the winner of draw s is excluded everywhere before draw s + 1
        V, I = all_ranks(*scorer.pick(s, seed))
        g = int(_winner(V, I))
        scorer.exclude(g)
        chosen.append(g)
    return X[torch.tensor(chosen, dtype=torch.long, device=X.device)]


def optimize_acqf_sharded(
    acq_function,
    bounds: Tensor,
    q: int,
    num_restarts: int,
    raw_samples: int,
    options: Optional[dict] = None,
    gen_candidates: Optional[Callable] = None,
    timings: Optional[dict] = None,
) -> Tuple[Tensor, Tensor]:
    """``optimize_acqf`` with raw samples and restarts sharded over the process group.

    Equivalent to the reference run with ``options['batch_limit'] = ceil(num_restarts / world_size)``.  Without
    ``options['seed']`` rank 0 draws the seed and broadcasts it, so every rank holds the same raw samples and makes the
    same host-side selection.  ``timings`` (optional dict) receives the wall time of each phase, device-synchronised.
    """
    from time import perf_counter

    from .optim import gen_candidates_scipy, initialize_q_batch_indices
    from .sampling import draw_sobol_samples

    options = dict(options or {})
    rank, ws = world()
    device = bounds.device

    def mark(name: str, t0: float) -> float:
        if timings is None:
            return t0
        if device.type == "cuda":
            torch.cuda.synchronize(device)
        t1 = perf_counter()
        timings[name] = timings.get(name, 0.0) + (t1 - t0)
        return t1

    t = perf_counter()
    seed = shared_seed(options.get("seed"), device)
    # the base samples of the q new points are drawn on a host thread while this one draws the raw samples
    pending = acq_function.prefetch_base_samples(q) if hasattr(acq_function, "prefetch_base_samples") else None
    # each rank draws only the rows of the raw samples it evaluates (the engine fast-forwarded to its block): the draw,
    # the box scaling and the host->device copy all shrink with the world size
    lo_r, hi_r = shard_range(raw_samples, rank, ws)
    X_loc = draw_sobol_samples(bounds=bounds.cpu(), n=raw_samples, q=q, seed=seed, rows=(lo_r, hi_r))
    if pending is not None:
        pending.join()
    t = mark("raw_sample_draw_s", t)
    Y_rnd = sharded_raw_sample_values(acq_function, X_loc, device, options.get("init_batch_limit"), total=raw_samples)
    t = mark("raw_sample_eval_and_allgather_s", t)
    # identical host RNG stream on every rank => identical selection everywhere; the selected rows are then fetched from
    # the ranks that drew them (one all-gather of the blocks, on the device under NCCL)
    gen = torch.random.get_rng_state()
    torch.manual_seed(seed or 0)
    idcs = initialize_q_batch_indices(Y=Y_rnd, n=num_restarts, **({"eta": options["eta"]} if "eta" in options else {}))
    torch.random.set_rng_state(gen)
    lo, hi = shard_range(num_restarts, rank, ws)
    if ws == 1:
        ics_local = X_loc[idcs[lo:hi]]
    else:
        X_dev = X_loc.to(device) if _backend() == "nccl" else X_loc
        ics_local = all_gather_blocks(X_dev, raw_samples)[idcs[lo:hi].to(X_dev.device)]
    t = mark("initial_condition_selection_s", t)
    gen_candidates = gen_candidates or gen_candidates_scipy
    gen_opts = {k: v for k, v in options.items() if k in ("maxiter", "method", "with_grad", "disp", "ftol", "gtol")}
    cands, vals = gen_candidates(ics_local.to(device), acq_function, lower_bounds=bounds[0], upper_bounds=bounds[1],
                                 options=gen_opts)  # fmt: skip
    t = mark("restart_optimisation_s", t)
    cand, val, _ = gather_best(vals, cands, lo)
    mark("best_gather_s", t)
    return cand, val


# ---- fully Bayesian fit: independent chains per rank (units that never interact) ------------------------------
def nuts_sample_sharded(make_potential, dim: int, chains_per_rank: int, num_samples: int, warmup_steps: int, device,
                        max_tree_depth: int = 6, seed: int = 0, init_radius: float = 2.0):
    """Every rank runs ``chains_per_rank`` lockstep NUTS chains on its own device (seed offset by the rank, so the chains of
    the job are distinct) and the draws are all-gathered: -> samples [num_samples x (world * chains_per_rank) x dim], chains
    in rank order, identical on every rank.  ``make_potential(num_chains)`` builds the rank-local potential.  One collective
    after the last draw; none inside the sampler."""
    from .fully_bayesian import nuts_sample

    rank, ws = world()
    potential = make_potential(chains_per_rank)
    gen = torch.Generator(device=device)
    gen.manual_seed(int(seed) + 7919 * rank)
    init = None
    for _ in range(20):  # redraw the starting points whose energy is not finite (init_to_uniform, radius 2)
        trial = 2.0 * init_radius * torch.rand(chains_per_rank, dim, generator=gen, device=device, dtype=torch.float64) - init_radius
        init = trial if init is None else torch.where(torch.isfinite(potential(init)[0]).unsqueeze(-1), init, trial)
        if bool(torch.isfinite(potential(init)[0]).all()):
            break
    samples, stats = nuts_sample(potential, init, num_samples, warmup_steps, max_tree_depth=max_tree_depth,
                                 seed=int(seed) + 7919 * rank + 1)  # fmt: skip
    if ws == 1:
        return samples, stats
    local = samples.transpose(0, 1).contiguous()  # chains x draws x dim: rank blocks concatenate along the chain axis
    out = [torch.empty_like(local) for _ in range(ws)]
    _all_gather(out, local)
    return torch.cat(out, dim=0).transpose(0, 1).contiguous(), stats


def fit_fully_bayesian_model_nuts_sharded(model, max_tree_depth: int = 6, warmup_steps: int = 512, num_samples: int = 256,
                                          thinning: int = 16, chains_per_rank: int = 1, seed: int = 0) -> None:
    """``fit_fully_bayesian_model_nuts`` with the chains spread over the ranks: each rank draws
    ``num_samples // (world * chains_per_rank)`` samples per chain after its own warm-up, every rank loads the same thinned
    ensemble (chain-major, rank order)."""
    from .fully_bayesian import SaasPotential

    _, ws = world()
    X = model.train_inputs[0]
    model.train()
    holder = {}

    def make(num_chains):
        holder["p"] = SaasPotential(X, model.train_targets, num_chains=num_chains)
        return holder["p"]

    per_chain = max(num_samples // (ws * chains_per_rank), 1)
    samples, stats = nuts_sample_sharded(make, X.shape[-1] + 4, chains_per_rank, per_chain, warmup_steps, X.device,
                                         max_tree_depth=max_tree_depth, seed=seed)  # fmt: skip
    kept = samples.transpose(0, 1).reshape(-1, samples.shape[-1])[::thinning]
    model.load_mcmc_samples(holder["p"].constrain(kept))
    model.eval()
    model.nuts_stats = {**stats, "potential_evaluations": holder["p"].evaluations}
