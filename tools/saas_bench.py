"""SAAS potential energy + gradient (the function NUTS evaluates) at the C4 shape n=100, d=30: batched over C
chains in one bo_mll_value_grad call, eager and as a CUDA-graph replay, against the CPU oracle; then a whole
fit_fully_bayesian_model_nuts.  Usage: python tools/saas_bench.py"""
import json
import sys
import time

import torch

sys.path.insert(0, ".")
import botorch_b200 as bb  # noqa: E402
from botorch_b200.fully_bayesian import SaasPotential  # noqa: E402
from oracle import saas as osaas  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
n, d = 100, 30
X = torch.rand(n, d, generator=g, dtype=torch.float64)
y = torch.sin(2 * torch.pi * X[:, 0]) + 0.1 * torch.randn(n, generator=g, dtype=torch.float64)
y = (y - y.mean()) / y.std()
out = {"n": n, "d": d, "potential": {}}
for C, graph in ((1, False), (1, True), (8, True), (64, True), (256, False), (256, True)):
    pot = SaasPotential(X.to(dev), y.to(dev), num_chains=C, graph=graph)
    U = (torch.rand(C, d + 4, generator=g, dtype=torch.float64) - 0.5).to(dev)
    for _ in range(5):
        pot(U)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    K = 30
    t0 = time.perf_counter()
    e0.record()
    for _ in range(K):
        pot(U)
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / K
    ms = e0.elapsed_time(e1) / K
    out["potential"][f"{C}{'_graph' if graph else '_eager'}"] = {"ms_per_call": ms, "wall_ms_per_call": wall * 1e3,
                                                                "chain_evals_per_s": C / (ms * 1e-3)}  # fmt: skip
U = torch.rand(16, d + 4, generator=g, dtype=torch.float64) - 0.5
osaas.potential_and_grad(U[:2], X, y)
t0 = time.perf_counter()
oe, og = osaas.potential_and_grad(U, X, y)
cpu = (time.perf_counter() - t0) / 16
pot = SaasPotential(X.to(dev), y.to(dev), num_chains=16)
e, gr = pot(U.to(dev))
out["cpu_oracle"] = {"ms_per_chain_eval": cpu * 1e3, "chain_evals_per_s": 1.0 / cpu, "cores": torch.get_num_threads(),
                     "energy_rel_err": float((e.cpu() - oe).abs().max() / oe.abs().max()),
                     "grad_rel_err": float((gr.cpu() - og).abs().max() / og.abs().max())}  # fmt: skip
for chains in (1, 8):
    model = bb.SaasFullyBayesianSingleTaskGP(X.to(dev), y.unsqueeze(-1).to(dev))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    bb.fit_fully_bayesian_model_nuts(model, warmup_steps=512, num_samples=256, thinning=16, num_chains=chains, seed=0)
    torch.cuda.synchronize()
    st = model.nuts_stats
    out[f"fit_nuts_chains{chains}"] = {
        "seconds": time.perf_counter() - t0, "potential_evaluations": st["potential_evaluations"],
        "accept_prob": float(st["accept_prob"].mean()), "tree_depth": float(st["tree_depth"].mean()),
        "divergent": int(st["divergent"].sum()), "retained": model.num_mcmc_samples,
        "median_lengthscale_dim0": float(model.median_lengthscale[0]),
        "median_lengthscale_others_min": float(model.median_lengthscale[1:].min())}  # fmt: skip
print(json.dumps(out))
