"""Generate the committed golden fixtures (run in the build container, where /root/reference exists).

    python tests/golden/make_golden.py

Part 1 imports the reference's OWN modules verbatim from /root/reference (through oracle/refstub.py, which
bypasses ``botorch/__init__.py`` because gpytorch is not installed) and records their outputs on seeded inputs:
  * botorch/utils/safe_math.py   log_fatplus / fatplus / fatmax / logmeanexp / logsumexp values AND gradients,
                                 including the +-inf cases of test/utils/test_safe_math.py:393-422
  * botorch/sampling/qmc.py, botorch/utils/sampling.py   Sobol-normal base samples, Sobol box samples
  * botorch/test_functions/synthetic.py                  Branin / Hartmann(6) / Ackley(20) values
  * botorch/optim/initializers.py cannot be imported (gpytorch); its argmax semantics are pinned by torch itself.
These pin the oracle (tests/test_oracle_golden.py) and, through it, the CUDA path.

Part 2 mints oracle outputs for small instances of the BASELINE.json configs (posterior mean/covariance, qLogNEI
value, gradient, raw-sample argmax).  GP posterior values have no reference-side fixture (gpytorch/linear_operator
are absent): for those rows parity is "unpinned" by the reference and pinned only against the oracle's two
independent formulations (see oracle/__init__.py).
"""
from __future__ import annotations

import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import refstub  # noqa: E402


def _nofat_part() -> dict:
    """fat = False pieces of the reference (utils/safe_math.py:226-273: log_softplus, smooth_amax), values and gradients,
    and the per-sample qLogNEI chain with them (acquisition/logei.py:122,532).  Own generator: the fixtures of
    `_reference_part` stay bit-identical.  -> tests/golden/reference_nofat.pt"""
    ref = refstub.load()
    sm = ref.safe_math
    g = torch.Generator().manual_seed(777)
    cases = {}
    x = torch.randn(64, 7, generator=g, dtype=torch.float64) * torch.logspace(-8, 2, 64, dtype=torch.float64)[:, None]
    for tau in (1e-6, 1e-2, 1.0):
        xg = x.clone().requires_grad_(True)
        v = sm.log_softplus(xg, tau=tau)
        (gr,) = torch.autograd.grad(v.sum(), xg)
        cases[f"log_softplus_tau{tau:g}"] = {"x": x, "value": v.detach(), "grad": gr}
    y = torch.randn(33, 5, 8, generator=g, dtype=torch.float64) * 3
    for tau in (1e-2, 1.0):
        yg = y.clone().requires_grad_(True)
        v = sm.smooth_amax(yg, dim=-1, tau=tau)
        (gr,) = torch.autograd.grad(v.sum(), yg)
        cases[f"smooth_amax_tau{tau:g}"] = {"x": y, "value": v.detach(), "grad": gr}
    inf = float("inf")
    z = torch.tensor([[1.0, -inf, 2.0, 0.5], [-inf, -inf, -inf, -inf], [inf, 1.0, inf, -3.0]], dtype=torch.float64)
    cases["smooth_amax_inf"] = {"x": z, "value": sm.smooth_amax(z, dim=-1, tau=1e-2)}
    obj = torch.randn(16, 6, 3, generator=g, dtype=torch.float64)
    best = torch.randn(16, 1, generator=g, dtype=torch.float64)
    og = obj.clone().requires_grad_(True)
    li = sm.log_softplus(og - best.unsqueeze(-1), tau=1e-6)
    acq = sm.logmeanexp(sm.smooth_amax(li, dim=-1, tau=1e-2), dim=0)
    (gr,) = torch.autograd.grad(acq.sum(), og)
    cases["qlognei_chain_nofat"] = {"obj": obj, "best_f": best, "value": acq.detach(), "grad": gr}
    return cases


def _reference_part() -> dict:
    ref = refstub.load()
    sm, out = ref.safe_math, {}
    g = torch.Generator().manual_seed(20240)

    # ---- safe_math: values + gradients ------------------------------------------------------------------------------
    x = torch.randn(64, 7, generator=g, dtype=torch.float64) * torch.logspace(-8, 2, 64, dtype=torch.float64)[:, None]
    cases = {}
    for tau in (1e-6, 1e-2, 1.0):
        xg = x.clone().requires_grad_(True)
        v = sm.log_fatplus(xg, tau=tau)
        (gr,) = torch.autograd.grad(v.sum(), xg)
        cases[f"log_fatplus_tau{tau:g}"] = {"x": x, "value": v.detach(), "grad": gr}
        v2 = sm.fatplus(x, tau=tau)
        cases[f"fatplus_tau{tau:g}"] = {"x": x, "value": v2}
    y = torch.randn(33, 5, 8, generator=g, dtype=torch.float64) * 3
    y[3, 1, 2] = y[3, 1, 5]  # a tie
    for tau in (1e-2, 1.0):
        yg = y.clone().requires_grad_(True)
        v = sm.fatmax(yg, dim=-1, tau=tau)
        (gr,) = torch.autograd.grad(v.sum(), yg)
        cases[f"fatmax_tau{tau:g}"] = {"x": y, "value": v.detach(), "grad": gr}
    yg = y.clone().requires_grad_(True)
    v = sm.logmeanexp(yg, dim=0)
    (gr,) = torch.autograd.grad(v.sum(), yg)
    cases["logmeanexp_dim0"] = {"x": y, "value": v.detach(), "grad": gr}
    # infinities (test_safe_math.py:393-422)
    inf = float("inf")
    z = torch.tensor([[1.0, -inf, 2.0, 0.5], [-inf, -inf, -inf, -inf], [inf, 1.0, inf, -3.0], [0.0, 0.0, 0.0, 0.0]],
                     dtype=torch.float64)  # fmt: skip
    for name, fn in (("fatmax_inf", lambda t: sm.fatmax(t, dim=-1, tau=1e-2)), ("logsumexp_inf", lambda t: sm.logsumexp(t, dim=-1)),
                     ("logmeanexp_inf", lambda t: sm.logmeanexp(t, dim=-1))):  # fmt: skip
        zg = z.clone().requires_grad_(True)
        v = fn(zg)
        (gr,) = torch.autograd.grad(v[torch.isfinite(v)].sum(), zg, allow_unused=True)
        cases[name] = {"x": z, "value": v.detach(), "grad": gr}
    # the full per-sample chain of qLogNEI (logei.py:509-534 + 122 + 129) on mock samples
    obj = torch.randn(16, 6, 3, generator=g, dtype=torch.float64)
    best = torch.randn(16, 1, generator=g, dtype=torch.float64)
    og = obj.clone().requires_grad_(True)
    li = sm.log_fatplus(og - best.unsqueeze(-1), tau=1e-6)
    acq = sm.logmeanexp(sm.fatmax(li, dim=-1, tau=1e-2), dim=0)
    (gr,) = torch.autograd.grad(acq.sum(), og)
    cases["qlognei_chain"] = {"obj": obj, "best_f": best, "value": acq.detach(), "grad": gr}
    # known answers of test/acquisition/test_logei.py:324-395 and :429-501: samples [0, 1] / zeros
    out["known_answers"] = {
        "single": {"baseline": torch.zeros(1, 1, dtype=torch.float64), "new": torch.ones(1, 1, 1, dtype=torch.float64), "exp_acq": 1.0},
        "batch": {"baseline": torch.zeros(1, 1, dtype=torch.float64),
                  "new": torch.tensor([[[1.0], [0.0]]], dtype=torch.float64), "exp_acq": [1.0, 0.0]},
    }  # fmt: skip
    out["safe_math"] = cases

    # ---- Sobol draws ----------------------------------------------------------------------------------------------------
    out["qmc"] = {
        f"normal_d{d}_n{n}_seed{seed}": ref.sampling.draw_sobol_normal_samples(d=d, n=n, dtype=torch.float64, seed=seed)
        for d, n, seed in ((3, 8, 0), (21, 16, 1234), (204, 32, 1234), (108, 4, 7))
    }
    bounds = torch.tensor([[0.0, -1.0, 2.0], [1.0, 1.0, 5.0]], dtype=torch.float64)
    out["qmc"]["box_n8_q2_seed0"] = ref.sampling.draw_sobol_samples(bounds=bounds, n=8, q=2, seed=0)
    out["qmc"]["box_n4_q1_batch3_seed5"] = ref.sampling.draw_sobol_samples(bounds=bounds, n=4, q=1, batch_shape=torch.Size([3]), seed=5)
    eng = ref.qmc.NormalQMCEngine(d=5, seed=11, inv_transform=False)
    out["qmc"]["boxmuller_d5_n6_seed11"] = eng.draw(6, dtype=torch.float64)

    # ---- synthetic functions ----------------------------------------------------------------------------------------------
    syn = ref.synthetic
    U = torch.rand(16, 20, generator=g, dtype=torch.float64)
    br, hm, ac = syn.Branin(), syn.Hartmann(dim=6), syn.Ackley(dim=20)
    Xb = torch.tensor(br._bounds, dtype=torch.float64).T
    Xb = Xb[0] + (Xb[1] - Xb[0]) * U[:, :2]
    Xa = -32.768 + 65.536 * U
    out["synthetic"] = {
        "branin": {"X": Xb, "f": br.evaluate_true(Xb)},
        "hartmann6": {"X": U[:, :6].clone(), "f": hm.evaluate_true(U[:, :6])},
        "ackley20": {"X": Xa, "f": ac.evaluate_true(Xa)},
    }
    return out


def _oracle_part() -> dict:
    from botorch_b200.synthetic import make_problem
    from tests.util import candidates, oracle_acq, oracle_gp

    out = {}
    for tag, cfg, kw, q, B in (("C1", "C1", {}, 1, 12), ("C2", "C2", {"n": 60}, 4, 10), ("C3", "C3", {"n": 96}, 8, 6),
                               ("C4", "C4", {"n": 40, "M": 3}, 4, 6)):  # fmt: skip
        p = make_problem(cfg, **kw)
        ogp = oracle_gp(p)
        oacq = oracle_acq(p, ogp)
        X = candidates(p, B, q=q)
        mean, cov = ogp.posterior(X)
        val, grad = oacq.value_and_grad(X, mode="faithful")
        raw = candidates(p, 128, q=q, seed=4242)
        with torch.no_grad():
            yraw = oacq.forward(raw, mode="faithful")
        out[tag] = {"cfg": cfg, "kw": kw, "q": q, "X": X, "mean": mean, "cov": cov, "acq": val, "grad": grad,
                    "raw_seed": 4242, "raw_argmax": int(torch.argmax(yraw)), "raw_max": float(yraw.max()),
                    "best_f": oacq.best_f.clone()}  # fmt: skip
    p = make_problem("C5", n=128)
    ogp = oracle_gp(p)
    Xc = candidates(p, 512, q=1, seed=1).squeeze(1)
    mean, cov = ogp.posterior(Xc.unsqueeze(1))
    out["C5"] = {"cfg": "C5", "kw": {"n": 128}, "X": Xc, "mean": mean.reshape(-1), "var": cov.reshape(-1),
                 "argmax_mean": int(torch.argmax(mean.reshape(-1)))}  # fmt: skip
    return out


def main() -> None:
    if "--nofat-only" in sys.argv:  # add the fat = False fixtures without touching the committed ones
        torch.save(_nofat_part(), os.path.join(HERE, "reference_nofat.pt"))
        print("reference_nofat.pt", os.path.getsize(os.path.join(HERE, "reference_nofat.pt")), "bytes")
        return
    torch.manual_seed(0)
    torch.save(_reference_part(), os.path.join(HERE, "reference_modules.pt"))
    torch.save(_oracle_part(), os.path.join(HERE, "oracle_configs.pt"))
    torch.save(_nofat_part(), os.path.join(HERE, "reference_nofat.pt"))
    for f in ("reference_modules.pt", "oracle_configs.pt", "reference_nofat.pt"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
