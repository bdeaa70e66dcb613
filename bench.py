#!/usr/bin/env python
"""Benchmark of the qLogNEI hot path (BASELINE.json metric) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C3]

One "step" = one fused forward + backward of qLogNEI over the restart batch of the config (C3: B = 512 restarts of
q = 8 points, n = 2000, d = 20, S = 256), i.e. exactly what one L-BFGS-B function evaluation of ``optimize_acqf``
costs (botorch/generation/gen.py:194-222).  N > 1 (torchrun, one rank per GPU): rank 0 computes the training-set
factors and broadcasts them (NCCL); every rank then steps its own block of B restarts -- weak scaling, no
data-path collective.  Prints ONE JSON line on rank 0.

``--impl reference`` times the CPU restatement of the reference algorithm (oracle/, faithful order of operations:
posterior over [X_baseline; X], L^-T GEMM covariance, cached-Cholesky sampling, autograd backward) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "qLogNEI fwd+bwd candidate evals/s"
UNIT = "candidate_evals/s"
# dram__bytes_read.sum + dram__bytes_write.sum of the four dgemm_dmma_kernel launches of one step, averaged per launch,
# from the `ncu --set full` capture of this command at the default workload (profiles/r01_ncu_full_step_v4.csv:
# 10.0 + 154.2 + 334.0 + 67.1 MB).  Algorithmic operand bytes of the same four launches: ~550 MB, i.e. no re-reads.
NCU_GEMM_TRAFFIC_BYTES_PER_LAUNCH = {("C3", 512): 141.3e6}  # superseded by profiles/r02_traffic.json (keyed by source digest)
FP64_DMMA_PEAK_TFLOPS = 37.07  # fallback only: tools/fp64_peaks.cu on this pool (profiles/r01_fp64_peaks.txt); the run measures its own


def _measure_fp64_peak(dev, stream):
    """DMMA.8x8x4 issue-rate microbenchmark of the library (bo_dmma_peak_launch), best of 5 ~25 ms launches, CUDA events."""
    import ctypes as C

    from botorch_b200 import _lib

    lib = _lib.lib()
    scratch = torch.zeros(1, dtype=torch.float64, device=dev)
    flops = C.c_double()
    best = 0.0
    for rep in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        _lib.check(lib.bo_dmma_peak_launch(50000, scratch.data_ptr(), C.byref(flops), _lib.stream_ptr(dev)))
        e1.record(stream)
        torch.cuda.synchronize(dev)
        if rep:
            best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


def _traffic_record(cfg: str, B: int):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture, valid only for the library
    sources it was taken from (profiles/r02_traffic.json holds the digest)."""
    from botorch_b200 import _lib

    path = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if not os.path.exists(path):
        return None, "no committed capture"
    rec = json.load(open(path))
    digest = _lib.lib().bo_source_digest().decode()
    key = f"{cfg}:{B}"
    if key not in rec.get("workloads", {}):
        return None, "no committed capture for this workload"
    if rec.get("source_digest") != digest:
        return None, f"the committed capture ({rec.get('capture')}) was taken from other library sources; re-capture"
    return rec["workloads"][key]["dram_bytes_per_launch"], f"{rec.get('capture')}: {rec['workloads'][key]['note']}"


def _args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C3")
    ap.add_argument("--restarts", type=int, default=None, help="override the restart batch per GPU")
    ap.add_argument("--cpu-sample", type=int, default=None, help="candidates in the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling sub-record")
    return ap.parse_args()


def _problem(cfg: str):
    from botorch_b200.synthetic import make_problem

    return make_problem(cfg)


def _restart_batch(p, B: int, seed: int = 0) -> torch.Tensor:
    """B x q x d synthetic restart batch (Sobol q-batches in the unit cube, seed != training-set seed)."""
    from botorch_b200.sampling import draw_sobol_samples

    bounds = torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double()
    return draw_sobol_samples(bounds, n=B, q=p.q, seed=777 + seed).contiguous()


def _algorithmic_flops_per_eval(n, n_b, d, q, S):
    """Forward and backward algorithmic flops of one candidate (t-batch) evaluation; DESIGN.md section 4."""
    fwd = 2 * (d + 2) * q * n + 2 * q * n + q * n * n + q * q * n + 2 * q * n * n_b + q * n_b * n_b + q * q * n_b \
        + q**3 / 3 + 2 * q * (n_b + q) * S + 30 * q * S  # fmt: skip
    bwd = 2 * q * n_b * S + q * n_b * n_b + 2 * q * n * n_b + q * n * n + 2 * q * q * (n + n_b) + 4 * q * n * d \
        + 40 * q * S  # fmt: skip
    return fwd, bwd


class _ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")  # fmt: skip

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(
                ["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=self.f, stderr=subprocess.DEVNULL)  # fmt: skip
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().strip().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))  # noqa: E702
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# -------------------------------------------------------------------------------------------------
# CPU leg (oracle): cpu_baseline of the default run and the whole --impl reference run
# -------------------------------------------------------------------------------------------------
def _cpu_setup(p):
    from oracle.gp import GPHyper, OracleGP
    from oracle.qlognei import OracleQLogNEI

    h = p.hypers
    gp = OracleGP(p.train_X, p.train_Y, GPHyper(p.kernel, h["lengthscale"], h["noise"], h["mean"], h.get("outputscale")),
                  standardize=p.standardize)  # fmt: skip
    return OracleQLogNEI(gp, p.train_X, num_samples=p.mc_samples, seed=1234)


def _cpu_step(acq, X, mode="faithful"):
    t0 = time.perf_counter()
    acq.value_and_grad(X, mode=mode)
    return time.perf_counter() - t0


def _cpu_baseline(p, chunk: int, budget_s: float = 10.0):
    """Oracle timed on the host cores on a bounded sample of the workload: chunks of ``chunk`` restarts (the reference's
    ``batch_limit``; its (n_b+q) x (2n+q) block per restart bounds the chunk by memory) until ~``budget_s`` seconds."""
    torch.set_num_threads(os.cpu_count() or 1)
    acq = _cpu_setup(p)
    X = _restart_batch(p, p.num_restarts, seed=1)
    _cpu_step(acq, X[:1])  # warm the caches (train Cholesky, L^-T) outside the timing
    chunk = min(chunk, X.shape[0])
    best, sample, lo = 0.0, 0, 0
    while best < budget_s and sample < 64 * X.shape[0]:  # small configs: several passes over the restarts fill the budget
        if lo + chunk > X.shape[0]:
            lo = 0
        best += _cpu_step(acq, X[lo : lo + chunk])
        sample += chunk
        lo += chunk
    # SURVEY.md 8(d) also asks for the minimal algorithm (only the q new rows, trsm order) on the same cores
    n_min = min(8 * chunk, p.num_restarts)
    Xm = _restart_batch(p, n_min, seed=1)
    _cpu_step(acq, Xm[:1], "minimal")
    t_min = _cpu_step(acq, Xm, "minimal")
    return {
        "minimal_algorithm": {"value": n_min / t_min, "unit": UNIT, "sample": f"{n_min} restarts, {t_min:.2f} s"},
        "value": sample / best,
        "unit": UNIT,
        "cores": torch.get_num_threads(),
        "kind": "port",
        "sample": f"{sample} restart evaluations in chunks of {chunk} (q x d = {p.q} x {p.d} each), fwd + autograd "
                  f"bwd, faithful reference order of operations, float64 torch CPU, {best:.2f} s",
    }


def _run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    p = _problem(args.config)
    torch.set_num_threads(os.cpu_count() or 1)
    sample = args.cpu_sample or (4 if p.n >= 1000 else 32)
    acq = _cpu_setup(p)
    X = _restart_batch(p, sample, seed=1)
    for _ in range(max(args.warmup, 1)):
        _cpu_step(acq, X[:1])
    t = [_cpu_step(acq, X) for _ in range(args.steps)]
    total = sum(t)
    value = sample * args.steps / total
    cores = torch.get_num_threads()
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": _config_dict(p, args.restarts or p.num_restarts),
        "note": f"CPU oracle restatement of the reference algorithm; each step = a bounded sample of {sample} of the "
                "restarts (fwd + autograd bwd), rate scaled per candidate",
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{sample} candidates per step, fwd + autograd bwd, faithful order of operations"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }  # fmt: skip
    print(json.dumps(line), flush=True)


def _mll_leg(p, dev, stream, cpu: bool, peak: float = FP64_DMMA_PEAK_TFLOPS):
    """bo_mll_value_grad at the workload's training set: CUDA-event time per evaluation, FP64 rate on its algorithmic
    flops (Cholesky n^3/3 + inverse factor n^3/3 + W = alpha alpha^T - Khat^-1 against the triangular factor n^3 +
    covariance / gradient contractions), and the CPU oracle (torch float64 + autograd) on the same problem."""
    import ctypes as C

    from botorch_b200 import _lib, engine

    y = ((p.train_Y - p.train_Y.mean()) / p.train_Y.std()).reshape(-1)
    h = p.hypers
    st = engine.MLLState(p.train_X.to(dev), y.to(dev), p.kernel, with_outputscale=h.get("outputscale") is not None)
    hy = [h["lengthscale"].reshape(-1).to(dev), h["noise"].reshape(()).to(dev), h["mean"].reshape(()).to(dev),
          None if h.get("outputscale") is None else h["outputscale"].reshape(()).to(dev)]  # fmt: skip
    st.evaluate(*hy)
    lib, gp = _lib.lib(), st.gp

    def raw():
        _lib.check(lib.bo_mll_value_grad(gp.ctx, C.byref(gp.c), st.work.data_ptr(), st.work.numel(), st.value.data_ptr(),
                                         st.grad.data_ptr(), gp.info.data_ptr(), _lib.stream_ptr(dev)))  # fmt: skip

    for _ in range(3):
        raw()
    torch.cuda.synchronize()
    reps = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        raw()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    n, d = p.n, p.d
    flops = n**3 * (1 / 3 + 1 / 3 + 1.0) + 4 * (d + 2) * n * n + 2 * (d + 1) * n * n
    leg = {"mll_evals_per_s": 1e3 / ms, "ms_per_eval": ms, "tflops_fp64": flops / (ms * 1e-3) / 1e12,
           "frac_of_fp64_peak": flops / (ms * 1e-3) / 1e12 / peak, "n": n, "d": d,
           "what": "bo_mll_value_grad: log N(y | m, K + s2 I) and its gradient w.r.t. lengthscales, noise, mean "
                   "(one fit_gpytorch_mll closure evaluation), hyperparameters resident in HBM"}  # fmt: skip
    if cpu:
        from oracle import mll as omll  # CPU baseline leg only

        hh = {"lengthscale": h["lengthscale"].reshape(-1), "noise": h["noise"].reshape(()), "mean": h["mean"].reshape(()),
              "outputscale": None if h.get("outputscale") is None else h["outputscale"].reshape(())}  # fmt: skip
        omll.value_and_grad(omll.log_marginal, p.train_X, y, p.kernel, hh)
        creps = 3 if n >= 1000 else 20
        t0 = time.perf_counter()
        for _ in range(creps):
            omll.value_and_grad(omll.log_marginal, p.train_X, y, p.kernel, hh)
        leg["cpu_oracle_ms_per_eval"] = 1e3 * (time.perf_counter() - t0) / creps
        leg["cpu_cores"] = os.cpu_count()
    return leg


def _constrained_leg(p, dev, B):
    """SURVEY 8f N2 at the workload's shape: qLogNEI over two outcomes (the workload's objective + one constrained outcome on
    the same inputs) with one smoothed constraint, forward + backward through the public API, CUDA-event timed."""
    import botorch_b200 as bb

    h = p.hypers
    g = torch.Generator().manual_seed(0)
    X = p.train_X
    y_con = (X[:, : min(4, p.d)].sum(-1) - 0.5 * min(4, p.d) + 0.05 * torch.randn(p.n, generator=g, dtype=torch.float64)).unsqueeze(-1)
    kw = {"lengthscale": h["lengthscale"], "noise": h["noise"], "mean_constant": h["mean"], "kernel": p.kernel}
    if h.get("outputscale") is not None:
        kw["outputscale"] = h["outputscale"]
    gps = [bb.SingleTaskGP(X.to(dev), y.to(dev), **kw) for y in (p.train_Y, y_con)]
    acqf = bb.qLogNoisyExpectedImprovement(
        bb.ModelListGP(*gps), X.to(dev), sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234),
        objective=bb.LinearMCObjective(torch.tensor([1.0, 0.0], dtype=torch.float64, device=dev)),
        constraints=[lambda Z: Z[..., 1] - 0.25], eta=1e-2)  # fmt: skip
    Xc = torch.rand(B, p.q, p.d, dtype=torch.float64, generator=g).to(dev)

    def step():
        Xg = Xc.detach().requires_grad_(True)
        v = acqf(Xg)
        return v, torch.autograd.grad(v.sum(), Xg)[0]

    for _ in range(3):
        v, gr = step()
    reps = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        v, gr = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return {"candidate_evals_per_s": B / (ms * 1e-3), "ms_per_step": ms, "outcomes": 2, "constraints": 1, "restarts": B,
            "finite": bool(torch.isfinite(v).all() and torch.isfinite(gr).all()),
            "what": "qLogNoisyExpectedImprovement(ModelListGP, constraints=[...]) forward + autograd backward: two fused "
                    "pipelines, bo_mc_reduce_fwd/_bwd, external sample adjoints (profiles/r01_constrained_c3.json)"}  # fmt: skip


def _saas_leg(dev, chains: int = 64):
    """the SAAS potential energy + gradient NUTS evaluates per leapfrog step, C4's training shape
    (n=100, d=30), `chains` chains per call as one CUDA-graph replay of bo_mll_value_grad + the prior terms."""
    from botorch_b200.fully_bayesian import SaasPotential

    g = torch.Generator().manual_seed(0)
    n, d = 100, 30
    X = torch.rand(n, d, generator=g, dtype=torch.float64)
    y = torch.sin(2 * torch.pi * X[:, 0]) + 0.1 * torch.randn(n, generator=g, dtype=torch.float64)
    pot = SaasPotential(X.to(dev), ((y - y.mean()) / y.std()).to(dev), num_chains=chains)
    U = (torch.rand(chains, d + 4, generator=g, dtype=torch.float64) - 0.5).to(dev)
    for _ in range(5):
        e, _ = pot(U)
    reps = 30
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        pot(U)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return {"chain_evals_per_s": chains / (ms * 1e-3), "ms_per_call": ms, "chains": chains, "n": n, "d": d,
            "finite": bool(torch.isfinite(e).all()),
            "what": "SaasPotential: potential energy + gradient of SaasPyroModel for all chains in one call "
                    "(profiles/r01_saas_potential_nuts.json)"}  # fmt: skip


def _config_dict(p, B):
    """`config` of the JSON line -- identical for both arms (`--impl ours` / `--impl reference`), so the driver can pair them."""
    n_models = int(p.hypers["lengthscale"].shape[0]) if p.hypers["lengthscale"].dim() == 2 else 1
    return {"workload": _workload_name(p, B), "kernel": p.kernel, "models": n_models,
            "l2": "inputs larger than L2: the n x n factor matrices plus the per-step workspace exceed the 126 MB L2 at the "
                  "default workload (C3: 192 MB + 219 MiB), no flush needed",
            "timed": "one fused forward + backward of qLogNEI over the restart batch per step (bo_qlognei_fwd + "
                     "bo_qlognei_bwd on the GPU arm, X resident in HBM; the oracle's value_and_grad on the reference arm)"}  # fmt: skip


def _workload_name(p, B):
    return (f"{p.name}: n={p.n} d={p.d} q={p.q} S={p.mc_samples} restarts/GPU={B} "
            f"(BASELINE.json configs[{'C1 C2 C3 C4 C5'.split().index(p.name[:2])}])")  # fmt: skip


# -------------------------------------------------------------------------------------------------
# GPU leg
# -------------------------------------------------------------------------------------------------
def _run_ours(args):
    import ctypes as C

    import torch.distributed as dist

    import botorch_b200 as bb
    from botorch_b200 import _lib, distributed as bdist

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if ws > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    p = _problem(args.config)
    B = args.restarts or p.num_restarts
    lib = _lib.lib()

    # ---- model + acquisition through the public API; factors built on rank 0 and broadcast --------------------
    h = p.hypers
    if p.kernel == "rbf":
        model = bb.SingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev), lengthscale=h["lengthscale"], noise=h["noise"],
                                mean_constant=h["mean"], outcome_transform="DEFAULT" if p.standardize else None)  # fmt: skip
    else:
        model = bb.SaasFullyBayesianSingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev))
        model.load_mcmc_samples(h)
    t_setup0 = time.perf_counter()
    sampler = bb.SobolQMCNormalSampler(sample_shape=torch.Size([p.mc_samples]), seed=1234)
    # rank 0 factors the training set and the baseline posterior; the other ranks receive them by NCCL broadcast
    acqf, bcast_bytes = bdist.make_qlognei(model, p.train_X.to(dev), sampler=sampler)
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t_setup0

    nei = acqf._nei
    X_host = _restart_batch(p, B, seed=rank).pin_memory()
    X = X_host.to(dev)
    acqf._set_sampler(p.q)
    Zq = acqf._Zq
    work = _lib.workspace(nei.workspace_bytes(B, p.q, True), dev)
    gin = -torch.ones(nei.gp.M, B, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev)

    def step():
        acq, info, _ = nei.forward(X, Zq, acqf.tau_relu, acqf.tau_max, with_grad=True, work=work)
        g = nei.backward(X, Zq, acqf.tau_relu, acqf.tau_max, gin, work)
        return acq, g

    def barrier():
        torch.cuda.synchronize()
        if ws > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    # ---- timed region: K steps, inputs resident in HBM, CUDA events on the launching stream ------------------------
    clocks = _ClockSampler(local) if rank == 0 else None
    time.sleep(0.3)
    lib.bo_launch_count_reset()
    # roofline of the DOMINANT kernel: the TMA-staged DMMA GEMM where the workload runs it (C3, C4, C5), else every GEMM launch
    prof_mode = 2 if args.config in ("C3", "C4") else 1
    lib.bo_prof_enable(prof_mode)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.profiler.start()  # `ncu --profile-from-start off` then sees exactly the timed region
    e0.record(stream)
    for _ in range(args.steps):
        acq, g = step()
    e1.record(stream)
    barrier()
    torch.cuda.profiler.stop()
    ms = e0.elapsed_time(e1)
    launches = int(lib.bo_launch_count())
    gemm_ms, gemm_n, gemm_fl = C.c_double(), C.c_long(), C.c_double()
    _lib.check(lib.bo_prof_read(C.byref(gemm_ms), C.byref(gemm_n), C.byref(gemm_fl)))
    lib.bo_prof_enable(0)
    assert bool(torch.isfinite(acq).all()) and bool(torch.isfinite(g).all())
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t)

    fp64_peak = _measure_fp64_peak(dev, stream) if rank == 0 else FP64_DMMA_PEAK_TFLOPS
    peak_note = ("DMMA.8x8x4 issue-rate microbenchmark run by this process right after the timed region (bo_dmma_peak_launch, "
                 "best of 5, CUDA events); MEASURED_PEAKS.json holds no FP64 figure")  # fmt: skip
    if not (20.0 < fp64_peak < 60.0):
        fp64_peak, peak_note = FP64_DMMA_PEAK_TFLOPS, "fallback constant (profiles/r01_fp64_peaks.txt): the in-run measurement failed"

    # ---- e2e: the user-facing call with HOST buffers (gen_candidates_scipy's f_np_wrapper, gen.py:194-222) -----------
    from botorch_b200.optim import HostEvaluator

    x_np = X_host.numpy().reshape(-1).copy()  # what SciPy hands to the objective: a plain float64 host array
    evaluate = HostEvaluator(acqf, X_host.shape, torch.empty(0, dtype=torch.float64, device=dev))

    def e2e_step():
        return evaluate(x_np)  # H2D of x, forward, autograd backward, D2H of (gradient, loss), one host wait

    for _ in range(3):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t)
    direct_route = evaluate.direct_calls > 0
    clk = clocks.stop() if clocks else None

    # ---- the whole candidate-generation loop (A4): SciPy L-BFGS-B on the host vs the device-resident optimiser -------------
    from botorch_b200.optim import gen_candidates_lbfgs, gen_candidates_scipy

    opt_iters = 15
    opt = {}
    for name, gen in (("scipy_lbfgsb_host", gen_candidates_scipy), ("lbfgs_device", gen_candidates_lbfgs)):
        if gen is gen_candidates_lbfgs:  # warm-up: capture the graphs of every batch size the compaction ladder visits
            gen(X, acqf, lower_bounds=0.0, upper_bounds=1.0, options={"maxiter": opt_iters})
        else:
            gen(X[: min(B, 8)], acqf, lower_bounds=0.0, upper_bounds=1.0, options={"maxiter": 2})
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        gen(X, acqf, lower_bounds=0.0, upper_bounds=1.0, options={"maxiter": opt_iters})
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if gen is gen_candidates_scipy:
            evals = gen_candidates_scipy.last_evaluator.calls + 1  # + the final no-grad evaluation
            cand_evals = B * evals
        else:  # the device optimiser drops stopped restarts from the batch: count restart evaluations
            stats = gen_candidates_lbfgs.last_stats
            evals, cand_evals = stats["rounds"] + 2, stats["restart_evaluations"] + B
        opt[name] = {"seconds": dt, "objective_evaluations": evals, "candidate_evals_per_s": cand_evals / dt,
                     "restarts_per_s": B / dt}

    # ---- posterior-only leg (second half of BASELINE.json's metric): marginal mean/var over a candidate stream --------
    Nc = 32768
    Xc = torch.rand(Nc, p.d, dtype=torch.float64, device=dev)
    pw = nei.gp.posterior_workspace(Nc)
    for _ in range(2):
        nei.gp.posterior_mean_var(Xc, work=pw)
    torch.cuda.synchronize()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record(stream)
    for _ in range(5):
        nei.gp.posterior_mean_var(Xc, work=pw)
    p1.record(stream)
    torch.cuda.synchronize()
    post_ms = p0.elapsed_time(p1) / 5
    post_flops = nei.gp.M * Nc * (2 * (p.d + 2) * p.n + 2 * p.n + p.n * p.n + 2 * p.n)

    # ---- strong scaling of ONE acquisition as configs[2] states it: 512 restarts + 32 768 raw samples in total, sharded -------
    # (tools/scaling.py: factor setup on rank 0 + one packed NCCL broadcast, raw-sample sweep + all-gather of Y, the rank's
    # block of restarts on the device optimiser, best-candidate gather; device-timed, max over ranks).  The single-GPU time
    # of the same job is measured in the same launch by rank 0 alone, so the line carries its own efficiency.
    strong = None
    if args.config == "C3" and not args.no_strong:
        from tools.scaling import _Solo, strong_c3

        # torchrun's OMP_NUM_THREADS=1 stays (os.cpu_count() is the host's, not the container's share: DESIGN.md section 7)
        bdist.warm_up(dev)
        world_group = dist.group.WORLD if ws > 1 else None
        r_1 = None
        if ws > 1:  # the single-GPU time of the same job first, on rank 0 alone; the other ranks SLEEP on the rendezvous
            # store until it is done (an NCCL barrier would spin one host core per idle rank next to the job being timed)
            store = dist.distributed_c10d._get_default_store()
            if rank == 0:
                with _Solo(bdist, True):
                    r_1 = strong_c3(dev, None, maxiter=20, factor_mode="recompute")
                store.set("bo_bench_solo_done", "1")
            else:
                store.wait(["bo_bench_solo_done"])
            barrier()
        r_n = strong_c3(dev, world_group, maxiter=20, factor_mode="broadcast")
        r_1 = r_n if r_1 is None else r_1
        strong = {"scaling": "strong", "what": "one optimize_acqf: 512 restarts + 32768 raw samples in total over N GPUs, "
                  "timed from the model's factorisation to the gathered best candidate (device L-BFGS, maxiter 20)",
                  "n_gpus": ws, "total_ms": r_n["total_ms"], "single_gpu_total_ms": r_1["total_ms"],
                  "speedup": r_1["total_ms"] / r_n["total_ms"], "efficiency": r_1["total_ms"] / r_n["total_ms"] / ws,
                  "setup_s": r_n["phases_rank0_s"]["factor_setup_and_broadcast_s"],
                  "single_gpu_setup_s": r_1["phases_rank0_s"]["factor_setup_and_broadcast_s"],
                  "phases_rank0_s": r_n["phases_rank0_s"], "single_gpu_phases_s": r_1["phases_rank0_s"],
                  "broadcast_bytes": r_n["broadcast_bytes"], "identical_on_all_ranks": r_n.get("identical_on_all_ranks"),
                  "best_value": r_n["best_value"], "single_gpu_best_value": r_1["best_value"],
                  "note": "the host-side Sobol engines of the reference's semantics (base samples in dimension n_b + q, raw "
                          "samples) do not shard: see DESIGN.md section 7"}  # fmt: skip

    if rank != 0:
        if ws > 1:
            dist.destroy_process_group()
        return
    # ---- hyperparameter side: exact MLL value + gradient per L-BFGS-B evaluation of fit_gpytorch_mll ----
    mll_leg = None
    if ws == 1 and nei.gp.M == 1:
        mll_leg = _mll_leg(p, dev, stream, cpu=not args.no_cpu_baseline, peak=fp64_peak)
    n_models = nei.gp.M
    fwd_f, bwd_f = _algorithmic_flops_per_eval(p.n, nei.n_b, p.d, p.q, p.mc_samples)
    survey_per_eval = 3 * fwd_f * n_models  # SURVEY.md 8(d) convention: fwd+bwd = 3 F_fwd of the trsm-based algorithm
    executed_per_eval = gemm_fl.value / (args.steps * B)  # GEMM flops this implementation runs
    value = ws * B * args.steps / (ms_max * 1e-3)
    traffic, traffic_note = _traffic_record(args.config, B)
    gemm_tflops = gemm_fl.value / (gemm_ms.value * 1e-3) / 1e12 if gemm_ms.value > 0 else 0.0
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ws, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": _config_dict(p, B),
        "workspace_mib": work.numel() / 2**20,
        "flops_per_eval": {"executed_gemm": executed_per_eval, "survey_8d_minimal_trsm_algorithm": survey_per_eval,
                           "note": "the combined-root path needs fewer flops than SURVEY 8(d)'s count; TFLOP/s figures "
                                   "use executed flops only"},
        "tflops_fp64_executed_whole_step": value * executed_per_eval / 1e12,
        "roofline": {
            "bound": "tensor", "achieved": gemm_tflops, "peak": fp64_peak, "unit": "TFLOP/s",
            "frac": gemm_tflops / fp64_peak, "traffic": traffic,
            "kernel": "dgemm_dmma_tma_kernel" if prof_mode == 2 else "dgemm_dmma_kernel (all GEMM launches)",
            "traffic_note": traffic_note, "peak_source": peak_note,
            "flops": "algorithmic: 2 x nonzeros(A) x N per launch (triangular operands count their nonzeros only)",
            "launches_per_step": gemm_n.value / args.steps, "avg_launch_ms": gemm_ms.value / max(gemm_n.value, 1),
            "share_of_step": gemm_ms.value / ms,
        },
        "e2e": {
            "value": ws * B * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": X_host.numel() * 8,
            "d2h_bytes_per_step": (X_host.numel() * 8 + B * 12) if direct_route else (X_host.numel() * 8 + 8),
            "api": "botorch_b200.optim.HostEvaluator (the objective gen_candidates_scipy hands to L-BFGS-B): host x in, "
                   + ("forward + reverse kernels with d loss / d acq = -1 into preallocated buffers (direct route), gradient + "
                      "per-restart values + failure flags out" if direct_route else
                      "qLogNoisyExpectedImprovement.forward + torch.autograd.grad, gradient + loss out"),
            "route": "direct" if direct_route else "autograd",
        },
        "gpu_launches": launches,
        "optimizer_loop": {"maxiter": opt_iters, "restarts": B, **opt,
                           "note": "wall clock of one gen_candidates call incl. its one-time setup (SciPy converts the bounds "
                                   "of all B*q*d variables in Python); rank 0 only"},
        "posterior": {"tflops_fp64": post_flops / (post_ms * 1e-3) / 1e12, "frac_of_fp64_peak":
                      post_flops / (post_ms * 1e-3) / 1e12 / fp64_peak, "candidates": Nc,
                      "ms": post_ms},  # fmt: skip
        "setup_s": setup_s, "broadcast_bytes": bcast_bytes, "clocks": clk,
    }  # fmt: skip
    if strong is not None:
        line["strong"] = strong
    if mll_leg is not None:
        line["mll"] = mll_leg
        line["constrained"] = _constrained_leg(p, dev, B)
        line["saas_potential"] = _saas_leg(dev)
    if not args.no_cpu_baseline and ws == 1:
        sample = args.cpu_sample or (4 if p.n >= 1000 else 32)
        line["cpu_baseline"] = _cpu_baseline(p, sample)
    print(json.dumps(line), flush=True)
    if ws > 1:
        dist.destroy_process_group()


def main():
    args = _args()
    if args.impl == "reference":
        _run_reference(args)
    else:
        _run_ours(args)


if __name__ == "__main__":
    main()
