"""Multi-GPU measurements of the sharded paths a user runs (BASELINE.json configs[2] and configs[4], plus the per-shape table).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
        tools/scaling.py [--what strong,c5,shapes] [--out gpurun_out/r02_scaling.json] [--c5-candidates 1000000]

ONE launch covers N = 1, 2, 4, ..., world size: for each N the first N ranks form a sub-group
(botorch_b200.distributed.use_group) and run the job while the others wait at a barrier, so every N is measured in the same
processes on the same box.  Every time is taken on the device (CUDA events on the launching stream, bracketed by a group
barrier + synchronize) and is the MAX over the group's ranks.

  strong  C3 as stated: ONE acquisition -- 512 restarts and 32 768 raw samples in total, split over N ranks.  Timed region =
          factor setup on rank 0 + broadcast (or per-rank recompute), raw-sample evaluation + all-gather of Y, the identical
          host-side selection, the restarts of the rank's block (device L-BFGS), best-candidate gather.
          (reference seam: botorch/optim/optimize.py:268-317, botorch/optim/initializers.py:416-423)
  c5      C5 as stated: exact GP n = 20 000, d = 12, marginal Thompson posterior over 1 000 000 candidates sharded over N ranks
          (generation/sampling.py:95-155), factor broadcast vs per-rank recompute, picks identical for every N.
  shapes  fwd+bwd candidate evals/s of each named shape (C1-C4), N ranks stepping their own restart block (weak scaling).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

FP64_PEAK = 37.07  # TFLOP/s, tools/fp64_peaks.cu on this pool (profiles/r01_fp64_peaks.txt)


def _group_sync(dev, group) -> None:
    torch.cuda.synchronize(dev)
    if group is not None and dist.get_world_size(group) > 1:
        dist.barrier(group=group)
    torch.cuda.synchronize(dev)


def _max_over(x: float, dev, group) -> float:
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    if group is not None and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t)


def _model(bb, p, dev):
    h = p.hypers
    if p.kernel == "rbf":
        return bb.SingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev), lengthscale=h["lengthscale"], noise=h["noise"],
                               mean_constant=h["mean"], outcome_transform="DEFAULT" if p.standardize else None)  # fmt: skip
    model = bb.SaasFullyBayesianSingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev))
    model.load_mcmc_samples(h)
    return model


# ----------------------------------------------------------------------------------------------------------- strong C3
def strong_c3(dev, group, maxiter: int = 20, factor_mode: str = "broadcast", reps: int = 3, num_restarts: int = 512,
              raw_samples: int = 32768, cfg: str = "C3", profile: bool = False):  # fmt: skip
    """One whole sharded acquisition, timed end to end on the device; returns the record of the last repetition (the
    first one warms the allocator pools, the pinned staging buffers and the group's NCCL channels)."""
    import botorch_b200 as bb
    from botorch_b200 import distributed as bd
    from botorch_b200.optim import gen_candidates_lbfgs
    from botorch_b200.synthetic import make_problem

    p = make_problem(cfg)
    bounds = torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double().to(dev)
    stream = torch.cuda.current_stream(dev)
    rank, ws = bd.world()
    rec = None
    for rep in range(reps):
        _group_sync(dev, group)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ph = {}
        prof = None
        if profile and rep == reps - 1 and rank == 0:
            import cProfile

            prof = cProfile.Profile()
            prof.enable()
        e0.record(stream)
        t0 = time.perf_counter()
        model = _model(bb, p, dev)
        sampler = bb.SobolQMCNormalSampler(sample_shape=torch.Size([p.mc_samples]), seed=1234)
        acqf, nbytes = bd.make_qlognei(model, p.train_X.to(dev), sampler=sampler, factor_mode=factor_mode)
        torch.cuda.synchronize(dev)
        ph["factor_setup_and_broadcast_s"] = time.perf_counter() - t0
        cand, val = bd.optimize_acqf_sharded(acqf, bounds, q=p.q, num_restarts=num_restarts, raw_samples=raw_samples,
                                             options={"seed": 7, "maxiter": maxiter}, gen_candidates=gen_candidates_lbfgs,
                                             timings=ph)  # fmt: skip
        e1.record(stream)
        torch.cuda.synchronize(dev)
        wall = time.perf_counter() - t0
        if prof is not None:
            import io
            import pstats

            prof.disable()
            buf = io.StringIO()
            pstats.Stats(prof, stream=buf).sort_stats("cumulative").print_stats(45)
            print(buf.getvalue(), file=sys.stderr, flush=True)
        ms = _max_over(e0.elapsed_time(e1), dev, group)
        stats = gen_candidates_lbfgs.last_stats
        evals = raw_samples + stats["restart_evaluations"] * ws  # candidate evaluations of the whole job (approx.: x ranks)
        rec = {
            "n_gpus": ws, "total_ms": ms, "wall_s_rank0": wall, "restarts_total": num_restarts, "raw_samples_total": raw_samples,
            "restarts_per_gpu": -(-num_restarts // ws), "maxiter": maxiter, "factor_mode": factor_mode,
            "broadcast_bytes": nbytes, "phases_rank0_s": {k: round(v, 6) for k, v in ph.items()},
            "lbfgs_rounds_rank0": stats["rounds"], "acquisitions_per_s": 1e3 / ms,
            "best_value": float(val), "best_candidate_checksum": float(cand.double().sum()),
            "candidate_evals_in_job_approx": int(evals),
        }  # fmt: skip
        del acqf, model
    # every rank must hold the same winner
    chk = torch.tensor([rec["best_value"], rec["best_candidate_checksum"]], dtype=torch.float64, device=dev)
    if ws > 1:
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
        rec["identical_on_all_ranks"] = bool(torch.equal(lo, hi))
    return rec


# ------------------------------------------------------------------------------------------------------------------ C5
def c5_sharded(dev, group, n: int = 20000, candidates: int = 1_000_000, chunk: int = 16384, factor_mode: str = "broadcast",
               num_samples: int = 2):  # fmt: skip
    """Marginal Thompson sampling over `candidates` points sharded across the group; setup and scoring timed apart."""
    import botorch_b200 as bb
    from botorch_b200 import distributed as bd
    from botorch_b200.synthetic import make_problem

    p = make_problem("C5", n=n)
    rank, ws = bd.world()
    stream = torch.cuda.current_stream(dev)
    _group_sync(dev, group)
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(stream)
    model = _model(bb, p, dev)
    nbytes = bd.replicate_model(model, factor_mode=factor_mode)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    ts = bb.MaxPosteriorSampling(model, chunk_size=chunk)
    # the candidate set is replicated (generated from its seed on every rank; only the rank's block goes to the device)
    Xc = torch.quasirandom.SobolEngine(p.d, scramble=True, seed=1).draw(candidates, dtype=torch.float64)
    lo, hi = bd.shard_range(candidates, rank, ws)
    Xd = Xc.to(dev)  # 96 MB at 1M x 12: replicated so that the winning row can be returned on every rank
    ts.marginal_posterior(Xd[lo : lo + min(chunk, hi - lo)])  # warm the workspace / kernels
    _group_sync(dev, group)
    e1b = torch.cuda.Event(enable_timing=True)
    e1b.record(stream)
    picks = bd.thompson_sample_sharded(ts, Xd, num_samples=num_samples, seed=3)
    e2.record(stream)
    torch.cuda.synchronize(dev)
    setup_ms = _max_over(e0.elapsed_time(e1), dev, group)
    score_ms = _max_over(e1b.elapsed_time(e2), dev, group)
    flops = candidates * (2.0 * (p.d + 2) * n + 2.0 * n + float(n) * n + 2.0 * n)
    idx = [int(torch.nonzero((Xc == pk.cpu()).all(-1))[0]) for pk in picks]
    st = model.gp_state
    return {
        "n_gpus": ws, "n": n, "d": p.d, "candidates_total": candidates, "candidates_per_gpu": hi - lo, "chunk": chunk,
        "factor_mode": factor_mode, "setup_ms": setup_ms, "broadcast_bytes": nbytes, "score_ms": score_ms,
        "candidates_per_s": candidates / (score_ms * 1e-3), "tflops_fp64": flops / (score_ms * 1e-3) / 1e12,
        "frac_of_fp64_peak_per_gpu": flops / (score_ms * 1e-3) / 1e12 / (FP64_PEAK * ws), "picked_global_indices": idx,
        "jitter": float(st.jitter.abs().max()), "draws": num_samples,
    }  # fmt: skip


def c5_cpu_baseline(n: int = 20000, candidates: int = 1024):
    """The oracle's marginal posterior (float64 torch CPU, all host cores) on a bounded sample of the C5 candidate stream:
    train factorisation timed apart from the per-candidate work (cross-covariance, triangular solve, column norms)."""
    from botorch_b200.synthetic import make_problem
    from oracle.gp import rbf

    torch.set_num_threads(os.cpu_count() or 1)
    p = make_problem("C5", n=n)
    h = p.hypers
    ls = h["lengthscale"].unsqueeze(-2)
    y = p.train_Y.squeeze(-1)
    targets = (y - y.mean()) / y.std()
    t0 = time.perf_counter()
    K = rbf(p.train_X, p.train_X, ls, x1_eq_x2=True)
    K.diagonal().add_(h["noise"])
    L = torch.linalg.cholesky(K)
    del K
    alpha = torch.cholesky_solve((targets - h["mean"]).unsqueeze(-1), L).squeeze(-1)
    setup_s = time.perf_counter() - t0
    Xc = torch.quasirandom.SobolEngine(p.d, scramble=True, seed=1).draw(candidates, dtype=torch.float64)
    t0 = time.perf_counter()
    Ktx = rbf(Xc, p.train_X, ls)
    mean = Ktx @ alpha
    V = torch.linalg.solve_triangular(L, Ktx.mT, upper=False)
    var = 1.0 - (V * V).sum(0)
    score_s = time.perf_counter() - t0
    assert bool(torch.isfinite(mean).all() and torch.isfinite(var).all())
    return {"kind": "port", "cores": torch.get_num_threads(), "n": n, "sample": f"{candidates} of the 1M candidates",
            "setup_s": setup_s, "score_s": score_s, "candidates_per_s": candidates / score_s}


# ------------------------------------------------------------------------------------------------- per-shape weak table
def shape_step(dev, group, cfg: str, steps: int = 20, warmup: int = 5):
    """bench.py's step (bo_qlognei_fwd + bo_qlognei_bwd over the config's restart batch, X resident) for one named shape."""
    import botorch_b200 as bb
    from botorch_b200 import _lib, distributed as bd
    from botorch_b200.sampling import draw_sobol_samples
    from botorch_b200.synthetic import make_problem

    p = make_problem(cfg)
    rank, ws = bd.world()
    lib = _lib.lib()
    model = _model(bb, p, dev)
    acqf, _ = bd.make_qlognei(model, p.train_X.to(dev), factor_mode="recompute",
                              sampler=bb.SobolQMCNormalSampler(sample_shape=torch.Size([p.mc_samples]), seed=1234))  # fmt: skip
    nei = acqf._nei
    B = p.num_restarts
    bounds = torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double()
    X = draw_sobol_samples(bounds, n=B, q=p.q, seed=777 + rank).contiguous().to(dev)
    acqf._set_sampler(p.q)
    Zq = acqf._Zq
    work = _lib.workspace(nei.workspace_bytes(B, p.q, True), dev)
    gin = -torch.ones(nei.gp.M, B, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev)

    def step():
        acq, info, _ = nei.forward(X, Zq, acqf.tau_relu, acqf.tau_max, with_grad=True, work=work)
        return acq, nei.backward(X, Zq, acqf.tau_relu, acqf.tau_max, gin, work)

    for _ in range(warmup):
        step()
    _group_sync(dev, group)
    lib.bo_prof_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        acq, g = step()
    e1.record(stream)
    _group_sync(dev, group)
    ms = _max_over(e0.elapsed_time(e1), dev, group)
    gemm_ms, gemm_n, gemm_fl = C.c_double(), C.c_long(), C.c_double()
    _lib.check(lib.bo_prof_read(C.byref(gemm_ms), C.byref(gemm_n), C.byref(gemm_fl)))
    lib.bo_prof_enable(0)
    assert bool(torch.isfinite(acq).all()) and bool(torch.isfinite(g).all())
    tf = gemm_fl.value / (gemm_ms.value * 1e-3) / 1e12 if gemm_ms.value > 0 else 0.0
    return {"config": cfg, "n_gpus": ws, "restarts_per_gpu": B, "models": nei.gp.M, "ms_per_step": ms / steps,
            "candidate_evals_per_s": ws * B * steps / (ms * 1e-3), "gemm_tflops_rank0": tf, "gemm_frac_of_fp64_peak": tf / FP64_PEAK,
            "gemm_share_of_step_rank0": gemm_ms.value / ms,
            "whole_step_tflops_per_gpu": gemm_fl.value / (ms * 1e-3) / 1e12}  # fmt: skip


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="strong,c5,shapes")
    ap.add_argument("--out", default=None)
    ap.add_argument("--maxiter", type=int, default=20)
    ap.add_argument("--c5-n", type=int, default=20000)
    ap.add_argument("--c5-candidates", type=int, default=1_000_000)
    ap.add_argument("--c5-cpu", action="store_true", help="also time the CPU oracle on a bounded C5 sample (rank 0)")
    ap.add_argument("--profile", action="store_true", help="cProfile the last strong-scaling repetition on rank 0")
    ap.add_argument("--sizes", default=None, help="comma list of group sizes (default: powers of two up to the world size)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from botorch_b200 import distributed as bd

    # torchrun exports OMP_NUM_THREADS=1 and it stays: os.cpu_count() reports the HOST's cores, not the container's share, and
    # giving each rank cpu_count / world OpenMP threads oversubscribed the box (the strong-scaling phases of the first 8-GPU
    # run of this tool were 3-5x slower on the host than the same function called from bench.py, which keeps the default).
    # The Sobol draws use their own small thread pool (botorch_b200/sampling.py).
    if ws > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    sizes = [int(s) for s in args.sizes.split(",")] if args.sizes else [n for n in (1, 2, 4, 8, 16) if n <= ws]
    groups = {}
    for n in sizes:  # new_group is collective over the whole job
        groups[n] = dist.new_group(ranks=list(range(n))) if ws > 1 and n > 1 else None
    what = args.what.split(",")
    out = {"world_size": ws, "gpu": torch.cuda.get_device_name(dev), "strong_c3": [], "c5": [], "shapes": [], "warmup_s": {}}
    for n in sizes:
        active = rank < n
        grp = groups[n]
        if active:
            ctx = bd.use_group(grp) if grp is not None else None
            if ctx is not None:
                ctx.__enter__()
            try:
                t0 = time.perf_counter()
                with _Solo(bd, n == 1):
                    bd.warm_up(dev)
                    out["warmup_s"][str(n)] = time.perf_counter() - t0
                    if "strong" in what:
                        for mode in ("broadcast", "recompute") if n > 1 else ("recompute",):
                            r = strong_c3(dev, grp, maxiter=args.maxiter, factor_mode=mode, profile=args.profile)
                            out["strong_c3"].append(r)
                            if rank == 0:
                                print("strong", json.dumps(r), file=sys.stderr, flush=True)
                    if "c5" in what:
                        for mode in ("broadcast", "recompute") if n > 1 else ("recompute",):
                            r = c5_sharded(dev, grp, n=args.c5_n, candidates=args.c5_candidates, factor_mode=mode)
                            out["c5"].append(r)
                            torch.cuda.empty_cache()
                            if rank == 0:
                                print("c5", json.dumps(r), file=sys.stderr, flush=True)
                    if "shapes" in what:
                        for cfg in ("C1", "C2", "C3", "C4"):
                            r = shape_step(dev, grp, cfg)
                            out["shapes"].append(r)
                            torch.cuda.empty_cache()
                            if rank == 0:
                                print("shape", json.dumps(r), file=sys.stderr, flush=True)
            finally:
                if ctx is not None:
                    ctx.__exit__(None, None, None)
        if ws > 1:  # ranks outside the group SLEEP on the rendezvous store while it works (an NCCL barrier spins a host core
            # per idle rank next to the job being timed); then everyone meets
            torch.cuda.synchronize(dev)
            store = dist.distributed_c10d._get_default_store()
            if rank == 0:
                store.set(f"bo_scaling_done_{n}", "1")
            else:
                store.wait([f"bo_scaling_done_{n}"])
            dist.barrier()
    if rank == 0 and "c5" in what and args.c5_cpu:
        out["c5_cpu_baseline"] = c5_cpu_baseline(args.c5_n)
        print("c5 cpu", json.dumps(out["c5_cpu_baseline"]), file=sys.stderr, flush=True)
    if ws > 1:
        dist.barrier()
    if rank == 0:
        t1 = {r["factor_mode"]: r["total_ms"] for r in out["strong_c3"] if r["n_gpus"] == 1}
        base = min(t1.values()) if t1 else None
        for r in out["strong_c3"]:
            if base:
                r["speedup_vs_1gpu"] = base / r["total_ms"]
                r["strong_efficiency"] = base / r["total_ms"] / r["n_gpus"]
        c1 = [r["score_ms"] for r in out["c5"] if r["n_gpus"] == 1]
        for r in out["c5"]:
            if c1:
                r["strong_efficiency_scoring"] = c1[0] / r["score_ms"] / r["n_gpus"]
        if out["c5"]:
            out["c5_picks_identical_for_every_n"] = len({tuple(r["picked_global_indices"]) for r in out["c5"]}) == 1
        text = json.dumps(out, indent=1)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, "w") as fh:
                fh.write(text)
        print(json.dumps(out))
    if ws > 1:
        dist.destroy_process_group()


class _Solo:
    """Inside a multi-rank launch, the N = 1 row runs on rank 0 alone: make botorch_b200.distributed see a world of one."""

    def __init__(self, bd, solo: bool) -> None:
        self.bd, self.solo, self.saved = bd, solo, None

    def __enter__(self):
        if self.solo:
            self.saved = self.bd.world
            self.bd.world = lambda: (0, 1)
        return self

    def __exit__(self, *exc) -> None:
        if self.solo:
            self.bd.world = self.saved


if __name__ == "__main__":
    main()
