"""The CUDA path (through the C ABI) against the 50-digit mpmath golden vectors of tests/golden/make_mpmath_golden.py:
posterior mean / covariance, baseline root, incumbents, qLogNEI value and gradient for n in {1, 3, 8}, q in {1, 2}, RBF-ARD
and scaled Matern-5/2, with and without Standardize, noise 1e-2 and 1e-4.  These vectors come from the published formulas
evaluated WITHOUT torch, so they pin the GPU path independently of the torch oracle.
Tolerance: BASELINE.json north_star -- 1e-9 relative in float64 (covariances relative to the prior variance)."""
import pytest
import torch

from tests.util import gp_state, mpmath_cases, oracle_gp, rel_err

pytestmark = pytest.mark.gpu

RTOL = 1e-9


@pytest.mark.parametrize("case,p", mpmath_cases(), ids=lambda v: v["name"] if isinstance(v, dict) else "")
def test_cuda_path_matches_mpmath_golden(cuda, case, p):
    from botorch_b200.engine import NEIState

    t = lambda v: torch.tensor(v, dtype=torch.float64)  # noqa: E731
    ogp = oracle_gp(p)  # only for the Standardize constants and standardised targets handed to the device state
    st = gp_state(p, ogp, cuda)
    X = t(case["X"])
    prior = case["y_std"] ** 2 * (case["outputscale"] or 1.0)
    mean, cov = st.posterior_joint(X.to(cuda))
    assert rel_err(mean[0], t(case["post_mean"])) < RTOL
    assert rel_err(cov[0], t(case["post_cov"]), scale=prior) < RTOL
    mv, vv = st.posterior_mean_var(X.reshape(-1, p.d).to(cuda))
    assert rel_err(mv.reshape(X.shape[:2]), t(case["post_mean"])) < RTOL
    assert rel_err(vv.reshape(X.shape[:2]), t(case["post_cov"]).diagonal(dim1=-1, dim2=-2), scale=prior) < RTOL
    nei = NEIState(st, p.train_X.to(cuda), t(case["Z_b"]).to(cuda))
    assert rel_err(nei.baseline_L[0], t(case["baseline_L"]), scale=prior**0.5) < RTOL
    assert rel_err(nei.best_f[0], t(case["best_f"])) < RTOL
    Zq = t(case["Z_q"]).to(cuda).contiguous()
    Xd = X.to(cuda)
    acq, info, work = nei.forward(Xd, Zq, 1e-6, 1e-2, with_grad=True)
    assert int(info.abs().sum()) == 0
    assert rel_err(acq[0], t(case["acq"]), scale=0.0) < RTOL
    g = nei.backward(Xd, Zq, 1e-6, 1e-2, torch.ones_like(acq), work)[0].cpu()
    gref = t(case["grad"])
    gerr = ((g - gref).flatten(1).abs().max(1).values / gref.flatten(1).abs().max(1).values).max()
    assert float(gerr) < RTOL
