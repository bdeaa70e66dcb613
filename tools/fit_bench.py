"""Time the exact-MLL value+gradient evaluation (bo_mll_value_grad) and a whole fit_gpytorch_mll at the BASELINE shapes,
next to the CPU oracle (torch float64 + autograd, all host cores).  Prints one JSON object.

usage: python tools/fit_bench.py [C2 C3 ...] [--no-fit]
"""
import ctypes as C
import json
import os
import sys
import time

import torch

sys.path.insert(0, ".")
from botorch_b200 import SingleTaskGP, _lib, engine  # noqa: E402
from botorch_b200.fit import ExactMarginalLogLikelihood, fit_gpytorch_mll  # noqa: E402
from botorch_b200.synthetic import make_problem  # noqa: E402
from oracle import mll as omll  # noqa: E402  (CPU baseline leg only)

dev = torch.device("cuda:0")
torch.set_num_threads(os.cpu_count())
cfgs = [a for a in sys.argv[1:] if not a.startswith("--")] or ["C2", "C3"]
out = {"cores": os.cpu_count(), "runs": {}}


def ev_time(fn, reps):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / reps


for name in cfgs:
    p = make_problem(name)
    n, d = p.train_X.shape
    X = p.train_X.to(dev)
    y = ((p.train_Y - p.train_Y.mean()) / p.train_Y.std()).reshape(-1)
    h = p.hypers
    st = engine.MLLState(X, y.to(dev), "rbf", with_outputscale=False)
    ls, nz, mu = h["lengthscale"].reshape(-1).to(dev), h["noise"].reshape(()).to(dev), h["mean"].reshape(()).to(dev)
    lib = _lib.lib()
    gp = st.gp

    def raw():  # the C-ABI call alone (hyperparameters resident)
        _lib.check(lib.bo_mll_value_grad(gp.ctx, C.byref(gp.c), st.work.data_ptr(), st.work.numel(), st.value.data_ptr(),
                                         st.grad.data_ptr(), gp.info.data_ptr(), _lib.stream_ptr(dev)))

    def setup_only():
        _lib.check(lib.bo_gp_setup(gp.ctx, C.byref(gp.c), st.work.data_ptr(), st.work.numel(), gp.info.data_ptr(),
                                   _lib.stream_ptr(dev)))

    st.evaluate(ls, nz, mu)
    reps = 20 if n <= 500 else 10
    ms_raw = ev_time(raw, reps)
    ms_setup = ev_time(setup_only, reps)
    # algorithmic flops: chol n^3/3 + inverse n^3/3 + W = [inv(L)^T | alpha] Bn (upper-triangular operand) n^3 + K 2(d+2)n^2 x2 + H 2(d+1)n^2
    flops = (n**3) * (1 / 3 + 1 / 3 + 1.0) + 4 * (d + 2) * n * n + 2 * (d + 1) * n * n
    r = {"n": n, "d": d, "mll_value_grad_ms": ms_raw, "of_which_factor_setup_ms": ms_setup,
         "mll_evals_per_s": 1e3 / ms_raw, "tflops_fp64": flops / (ms_raw * 1e-3) / 1e12}
    # end to end through the public closure: host x in, value + gradient out (what L-BFGS-B sees)
    model = SingleTaskGP(X, y.unsqueeze(-1).to(dev), outcome_transform=None)
    mll = ExactMarginalLogLikelihood(model.likelihood, model)
    from botorch_b200.fit import NdarrayOptimizationClosure, get_loss_closure_with_grads, get_parameters

    params = get_parameters(mll, requires_grad=True)
    nd = NdarrayOptimizationClosure(get_loss_closure_with_grads(mll, params), params)
    x0 = nd.state
    for _ in range(3):
        nd(x0)
    t0 = time.perf_counter()
    for _ in range(reps):
        nd(x0)
    r["e2e_closure_ms"] = 1e3 * (time.perf_counter() - t0) / reps
    # CPU oracle: same objective, torch float64 autograd
    Xc, yc = p.train_X, y
    hh = {"lengthscale": h["lengthscale"].reshape(-1), "noise": h["noise"].reshape(()), "mean": h["mean"].reshape(()), "outputscale": None}
    omll.value_and_grad(omll.exact_mll, Xc, yc, "rbf", hh)
    creps = 3 if n >= 1000 else 20
    t0 = time.perf_counter()
    for _ in range(creps):
        omll.value_and_grad(omll.exact_mll, Xc, yc, "rbf", hh)
    r["cpu_oracle_ms"] = 1e3 * (time.perf_counter() - t0) / creps
    if "--no-fit" not in sys.argv:
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fit_gpytorch_mll(mll, optimizer_kwargs={"options": {"maxiter": 50}})
        torch.cuda.synchronize()
        r["fit_s"] = time.perf_counter() - t0
        r["fit_evaluations"] = mll.evaluations
        r["fit_final_mll"] = float(mll.value_and_grads()[0])
    out["runs"][name] = r
print(json.dumps(out))
