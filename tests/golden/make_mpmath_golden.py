"""Independent 50-digit golden vectors for the GP-posterior / cached-Cholesky / qLogNEI algebra (TEST INFRASTRUCTURE).

    python tests/golden/make_mpmath_golden.py        # writes tests/golden/mpmath_gp.json (committed)

Why: the arithmetic of `model.posterior` lives in gpytorch 1.12 / linear_operator 0.5.2, which are not vendored under
/root/reference and cannot be installed here, and the reference's tests hold no golden vectors for posterior values
(SURVEY.md 8c).  `oracle/gp.py` restates those packages' algorithms in torch float64; this script pins that restatement
(and the CUDA path) against the PUBLISHED FORMULAS evaluated WITHOUT torch, in 50-digit `mpmath` arithmetic, so that a
shared misreading (distance centring, the Matern clamp, the prediction caches, the cached root update) cannot hide
behind torch.  Nothing below imports torch, the oracle or the product for the arithmetic; torch's `SobolEngine` is used
only to draw the base samples, which are INPUTS here (bit-pinned against the reference's own sampler module elsewhere:
tests/test_oracle_golden.py::test_sobol_draws_match_reference).

Formulas (reference call sites in /root/reference/botorch):
  * kernels: RBF-ARD `exp(-1/2 sum_t ((x_t-y_t)/l_t)^2)` (models/utils/gpytorch_modules.py:100-127);
    ScaleKernel(Matern-5/2 ARD) `os (1 + sqrt5 r + 5/3 r^2) exp(-sqrt5 r)` (models/fully_bayesian.py:81-85,275-282)
  * exact GP posterior behind `mvn = self(X)` (models/gpytorch.py:446): `mu = m + k*^T (K + s2 I)^-1 (y - m)`,
    `Sigma = k** - k*^T (K + s2 I)^-1 k*`
  * `Standardize` (models/transforms/outcome.py:289-307,373-447): `y -> (y - mean) / std` (unbiased std, 1 if n == 1),
    posterior `mu -> mean + std mu`, `Sigma -> std^2 Sigma`
  * cached-root sampling (utils/low_rank.py:118-162, acquisition/cached_cholesky.py:94-165, acquisition/logei.py:391-481):
    `L_b = chol(Sigma_bb)`, `bl_chol = Sigma_qb L_b^-T`, `br_chol = chol(Sigma_qq - bl_chol bl_chol^T)`,
    `f = mu_q + [bl_chol, br_chol] [z_b; z_q]`, `best_f[s] = max_j (mu_b + L_b z_b[s])_j`
  * reductions (utils/safe_math.py:293-352,209-223; acquisition/logei.py:509-534): `log_fatplus` with torch's softplus
    (identity above the threshold 20 -- that is what the reference executes), `fatmax` (alpha = 2), `logmeanexp`
  * gradient: `mpmath.diff` (high-order finite differences at 50 digits, ~25 correct digits)
"""
from __future__ import annotations

import itertools
import json
import math
import os
import sys

import mpmath as mp

HERE = os.path.dirname(os.path.abspath(__file__))
mp.mp.dps = 50

TAU_RELU, TAU_MAX = mp.mpf("1e-6"), mp.mpf("1e-2")  # acquisition/logei.py:66-67
S = 8  # MC samples of the fixture cases
SEED = 4321


# ------------------------------------------------------------------------------------------------- linear algebra (mp)
def chol(A):
    n = len(A)
    L = [[mp.mpf(0)] * n for _ in range(n)]
    for j in range(n):
        s = A[j][j] - sum(L[j][k] ** 2 for k in range(j))
        assert s > 0, "not positive definite in exact arithmetic"
        L[j][j] = mp.sqrt(s)
        for i in range(j + 1, n):
            L[i][j] = (A[i][j] - sum(L[i][k] * L[j][k] for k in range(j))) / L[j][j]
    return L


def solve_lower(L, b):  # L x = b
    n = len(L)
    x = [mp.mpf(0)] * n
    for i in range(n):
        x[i] = (b[i] - sum(L[i][k] * x[k] for k in range(i))) / L[i][i]
    return x


def solve_upper_t(L, b):  # L^T x = b
    n = len(L)
    x = [mp.mpf(0)] * n
    for i in reversed(range(n)):
        x[i] = (b[i] - sum(L[k][i] * x[k] for k in range(i + 1, n))) / L[i][i]
    return x


# ------------------------------------------------------------------------------------------------------------- kernels
def kernel(kind, x, y, ls, os_):
    r2 = sum(((a - b) / l) ** 2 for a, b, l in zip(x, y, ls))
    if kind == "rbf":
        return mp.exp(-r2 / 2)
    r = mp.sqrt(max(r2, mp.mpf("1e-30")))  # gpytorch MaternKernel: dist = sqrt(clamp_min(sq_dist, 1e-30))
    s5 = mp.sqrt(5)
    return os_ * (1 + s5 * r + mp.mpf(5) / 3 * r2_clamped(r)) * mp.exp(-s5 * r)


def r2_clamped(r):
    return r * r


# ---------------------------------------------------------------------------------------------------------- reductions
def softplus_torch(u):
    """torch.nn.functional.softplus(beta=1, threshold=20): identity above the threshold."""
    return u if u > 20 else mp.log1p(mp.exp(u))


def log_fatplus(x, tau):
    u = x / tau
    return mp.log(tau * (softplus_torch(u) + mp.mpf("0.1") / (1 + u * u)))


def fatmax(vals, tau):
    M = max(vals)
    acc = mp.mpf(0)
    for v in vals:
        u = (M - v) / tau
        acc += 2 / (2 + 2 * u + u * u)  # _pareto with alpha = 2
    return M + tau * mp.log(acc)


def logmeanexp(vals):
    M = max(vals)
    return M + mp.log(sum(mp.exp(v - M) for v in vals)) - mp.log(len(vals))


# --------------------------------------------------------------------------------------------------------------- model
class MpGP:
    def __init__(self, kind, train_X, train_Y, ls, os_, noise, mean_const, standardize):
        self.kind, self.ls, self.os, self.noise, self.m = kind, ls, os_, noise, mean_const
        self.X = train_X
        n = len(train_X)
        if standardize:
            self.y_mean = sum(train_Y) / n
            if n > 1:
                sd = mp.sqrt(sum((y - self.y_mean) ** 2 for y in train_Y) / (n - 1))
                self.y_std = sd if sd >= mp.mpf("1e-8") else mp.mpf(1)
            else:
                self.y_std = mp.mpf(1)
        else:
            self.y_mean, self.y_std = mp.mpf(0), mp.mpf(1)
        self.t = [(y - self.y_mean) / self.y_std for y in train_Y]
        K = [[kernel(kind, a, b, ls, os_) for b in train_X] for a in train_X]
        for i in range(n):
            K[i][i] = (os_ if kind != "rbf" else mp.mpf(1)) + noise  # k(x, x) exactly (x1_eq_x2 path) + s2
        self.L = chol(K)
        self.alpha = solve_upper_t(self.L, solve_lower(self.L, [t - mean_const for t in self.t]))

    def posterior(self, pts):
        """mean [p], cov [p][p] on the outcome scale for a list of points."""
        kx = [[kernel(self.kind, x, a, self.ls, self.os) for a in self.X] for x in pts]
        V = [solve_lower(self.L, row) for row in kx]  # rows: L^-1 k(X, x)
        mean = [self.y_mean + self.y_std * (self.m + sum(k * a for k, a in zip(row, self.alpha))) for row in kx]
        p = len(pts)
        cov = [[None] * p for _ in range(p)]
        for i in range(p):
            for j in range(p):
                kij = kernel(self.kind, pts[i], pts[j], self.ls, self.os)
                cov[i][j] = self.y_std**2 * (kij - sum(a * b for a, b in zip(V[i], V[j])))
        return mean, cov


class MpQLogNEI:
    def __init__(self, gp: MpGP, Z_b):
        self.gp, self.Z_b = gp, Z_b
        self.nb = len(gp.X)
        mean_b, cov_b = gp.posterior(gp.X)
        self.mean_b, self.cov_b = mean_b, cov_b
        self.L_b = chol(cov_b)
        self.base = [[mean_b[i] + sum(self.L_b[i][k] * z[k] for k in range(i + 1)) for i in range(self.nb)] for z in Z_b]
        self.best_f = [max(row) for row in self.base]

    def samples(self, Xq, Z_q):
        nb, q = self.nb, len(Xq)
        mean, cov = self.gp.posterior(list(self.gp.X) + list(Xq))
        bl = [[cov[nb + i][j] for j in range(nb)] for i in range(q)]
        br = [[cov[nb + i][nb + j] for j in range(q)] for i in range(q)]
        bl_chol = [solve_lower(self.L_b, row) for row in bl]  # rows of Sigma_qb L_b^-T
        schur = [[br[i][j] - sum(a * b for a, b in zip(bl_chol[i], bl_chol[j])) for j in range(q)] for i in range(q)]
        br_chol = chol(schur)
        out = []
        for zb, zq in zip(self.Z_b, Z_q):
            out.append([
                mean[nb + i] + sum(a * z for a, z in zip(bl_chol[i], zb)) + sum(br_chol[i][k] * zq[k] for k in range(i + 1))
                for i in range(q)
            ])
        return out, mean[nb:], br

    def value(self, Xq, Z_q):
        f, _, _ = self.samples(Xq, Z_q)
        per_sample = [fatmax([log_fatplus(v - bf, TAU_RELU) for v in row], TAU_MAX) for row, bf in zip(f, self.best_f)]
        return logmeanexp(per_sample)


# --------------------------------------------------------------------------------------------------------------- inputs
def _inputs(n, q, d, B, rng):
    train_X = [[rng.random() for _ in range(d)] for _ in range(n)]
    train_Y = [math.sin(6.0 * x[0]) + 0.5 * math.cos(3.0 * x[1]) + 0.3 * x[2] + 0.05 * rng.gauss(0, 1) + 2.5 for x in train_X]
    X = [[[rng.random() for _ in range(d)] for _ in range(q)] for _ in range(B)]
    return train_X, train_Y, X


def _base_samples(n_b, q):
    """The draws qLogNEI ends up holding (sampling/normal.py:68-131,178-209): inputs of the mp computation."""
    import torch
    from torch.quasirandom import SobolEngine

    def normal(dim):
        u = SobolEngine(dimension=dim, scramble=True, seed=SEED).draw(S, dtype=torch.float64)
        v = 0.5 + (1 - torch.finfo(torch.float64).eps) * (u - 0.5)
        return torch.erfinv(2 * v - 1) * math.sqrt(2)

    return normal(n_b).tolist(), normal(n_b + q)[:, n_b:].tolist()


def _f(x):
    return float(x)


def main():
    import random

    d, B = 3, 2
    cases = []
    grid = itertools.product(("rbf", "matern52"), (1, 3, 8), (1, 2), (True, False), (1e-2, 1e-4))
    for idx, (kind, n, q, standardize, noise) in enumerate(grid):
        rng = random.Random(1000 + idx)
        train_X, train_Y, X = _inputs(n, q, d, B, rng)
        ls = [0.35, 0.6, 1.4]
        os_ = 1.7 if kind == "matern52" else None
        mean_const = 0.15
        Z_b, Z_q = _base_samples(n, q)
        to_mp = lambda a: [mp.mpf(v) for v in a]  # noqa: E731  (float64 -> mpf is exact)
        gp = MpGP(kind, [to_mp(x) for x in train_X], to_mp(train_Y), to_mp(ls), mp.mpf(os_) if os_ else mp.mpf(1),
                  mp.mpf(noise), mp.mpf(mean_const), standardize)  # fmt: skip
        acq = MpQLogNEI(gp, [to_mp(z) for z in Z_b])
        zq = [to_mp(z) for z in Z_q]
        post_mean, post_cov, samples, values, grads = [], [], [], [], []
        for Xb in X:
            Xmp = [to_mp(x) for x in Xb]
            m, c = gp.posterior(Xmp)
            post_mean.append([_f(v) for v in m])
            post_cov.append([[_f(v) for v in row] for row in c])
            f, _, _ = acq.samples(Xmp, zq)
            samples.append([[_f(v) for v in row] for row in f])
            values.append(_f(acq.value(Xmp, zq)))

            def flat_value(*args):
                pts = [list(args[i * d : (i + 1) * d]) for i in range(q)]
                return acq.value(pts, zq)

            flat = [v for x in Xmp for v in x]
            g = []
            for k in range(q * d):
                order = tuple(1 if i == k else 0 for i in range(q * d))
                g.append(_f(mp.diff(flat_value, tuple(flat), order)))
            grads.append([g[i * d : (i + 1) * d] for i in range(q)])
        cases.append({
            "name": f"{kind}-n{n}-q{q}-{'std' if standardize else 'raw'}-noise{noise:g}",
            "kernel": kind, "n": n, "q": q, "d": d, "standardize": standardize, "noise": noise, "lengthscale": ls,
            "outputscale": os_, "mean_const": mean_const, "seed": SEED, "S": S, "train_X": train_X, "train_Y": train_Y,
            "X": X, "Z_b": Z_b, "Z_q": Z_q,
            "y_mean": _f(gp.y_mean), "y_std": _f(gp.y_std),
            "post_mean": post_mean, "post_cov": post_cov,
            "baseline_mean": [_f(v) for v in acq.mean_b],
            "baseline_L": [[_f(v) for v in row] for row in acq.L_b],
            "best_f": [_f(v) for v in acq.best_f],
            "samples": samples,  # [B][S][q]
            "acq": values, "grad": grads,
        })  # fmt: skip
        print(cases[-1]["name"], values, file=sys.stderr)
    out = {"meta": {"generator": "tests/golden/make_mpmath_golden.py", "mpmath": mp.__version__, "dps": mp.mp.dps,
                    "tau_relu": float(TAU_RELU), "tau_max": float(TAU_MAX)}, "cases": cases}  # fmt: skip
    with open(os.path.join(HERE, "mpmath_gp.json"), "w") as fh:
        json.dump(out, fh)
    print(f"wrote {len(cases)} cases", file=sys.stderr)


if __name__ == "__main__":
    main()
