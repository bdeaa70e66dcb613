"""Shared helpers for the parity tests: build the oracle and the CUDA state from one Problem."""
from __future__ import annotations

import torch

from botorch_b200.synthetic import Problem, make_problem
from oracle import qmc
from oracle.gp import GPHyper, OracleGP
from oracle.qlognei import OracleQLogNEI

CAND_SEED = 777  # != the training-set Sobol seed: identical seeds make candidates coincide with train points


def oracle_gp(p: Problem) -> OracleGP:
    h = GPHyper(p.kernel, p.hypers["lengthscale"], p.hypers["noise"], p.hypers["mean"], p.hypers.get("outputscale"))
    return OracleGP(p.train_X, p.train_Y, h, standardize=p.standardize)


def oracle_acq(p: Problem, gp: OracleGP = None, seed: int = 1234, X_baseline=None) -> OracleQLogNEI:
    gp = gp or oracle_gp(p)
    Xb = p.train_X if X_baseline is None else X_baseline
    return OracleQLogNEI(gp, Xb, num_samples=p.mc_samples, seed=seed)


def candidates(p: Problem, B: int, q: int = None, seed: int = CAND_SEED) -> torch.Tensor:
    q = q or p.q
    bounds = torch.stack([torch.zeros(p.d), torch.ones(p.d)]).double()
    return qmc.sobol_box(bounds, B, q, seed=seed)


def gp_state(p: Problem, ogp: OracleGP, device):
    from botorch_b200.engine import GPState

    h = p.hypers
    return GPState(
        p.train_X.to(device),
        ogp.train_targets.to(device),
        p.kernel,
        h["lengthscale"].to(device),
        h["noise"].to(device),
        h["mean"].to(device),
        None if h.get("outputscale") is None else h["outputscale"].to(device),
        y_mean=float(ogp.y_mean),
        y_std=float(ogp.y_std),
    )


def rel_err(a: torch.Tensor, b: torch.Tensor, scale=None) -> float:
    """max |a-b| / max(|b|, scale) elementwise; scale defaults to max|b| (matrix-relative)."""
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    s = b.abs().max() if scale is None else torch.as_tensor(scale)
    return float(((a - b).abs() / torch.maximum(b.abs(), s.expand_as(b))).max())


def mpmath_cases():
    """The 50-digit golden vectors of tests/golden/make_mpmath_golden.py as (case dict, Problem) pairs."""
    import json
    import os

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "mpmath_gp.json")
    with open(path) as fh:
        data = json.load(fh)
    out = []
    for c in data["cases"]:
        t = lambda v: torch.tensor(v, dtype=torch.float64)  # noqa: E731
        hypers = {"lengthscale": t(c["lengthscale"]), "noise": t(c["noise"]), "mean": t(c["mean_const"])}
        if c["outputscale"] is not None:
            hypers["outputscale"] = t(c["outputscale"])
        p = Problem(c["name"], t(c["train_X"]), t(c["train_Y"]).unsqueeze(-1), c["kernel"], hypers, c["q"], c["S"], 0, 0,
                    standardize=c["standardize"])  # fmt: skip
        out.append((c, p))
    return out
