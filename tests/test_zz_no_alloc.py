"""bo_kernel_matrix_static (the entry point that allocates nothing: its scratch is a file-scope device array) against
bo_kernel_matrix (caller-owned workspace), through the C ABI.  Collected last on purpose (file name):
a failure here must not hide any other test.  Passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

from tests.util import candidates, gp_state, oracle_gp, rel_err

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_DOUBLES = 64 * 1024  # BO_STATIC_SCRATCH_DOUBLES in include/botorch_b200.h


def _static(st, Xc):
    from botorch_b200 import _lib

    Nc = Xc.shape[0]
    out = torch.full((st.M, st.n, _lib.pad8(Nc)), float("nan"), dtype=torch.float64, device=st.device)
    code = _lib.lib().bo_kernel_matrix_static(st.ctx, C.byref(st.c), Nc, Xc.data_ptr(), out.data_ptr(),
                                              _lib.stream_ptr(st.device))  # fmt: skip
    return code, out[:, :, :Nc]


@pytest.mark.parametrize("cfg,kw", [("C2", {}), ("C4", {"M": 3})])
def test_static_scratch_entry_point_matches_the_workspace_one_twice_in_a_row(cuda, cfg, kw):
    from botorch_b200 import _lib
    from botorch_b200.synthetic import make_problem

    p = make_problem(cfg, **kw)
    ogp = oracle_gp(p)
    st = gp_state(p, ogp, cuda)
    # two calls in a row with different candidate sets and widths: the second must not see the first one's scratch
    for B, seed in ((61, 11), (7, 12)):
        Xc = candidates(p, B, seed=seed).reshape(-1, p.d).to(cuda).contiguous()
        ref = st.kernel_matrix(Xc)
        code, got = _static(st, Xc)
        _lib.check(code)
        torch.cuda.synchronize()
        assert bool(torch.isfinite(got).all())
        assert rel_err(got, ref) < 1e-13  # the tolerance of test_gpu_blocks.py; same kernels, so the bits agree
        assert torch.equal(got, ref)
        Kref = ogp._kernel(Xc.cpu().unsqueeze(-3) if ogp.hyper.batch_shape else Xc.cpu(), p.train_X).mT
        assert rel_err(got.reshape(Kref.shape) if not ogp.hyper.batch_shape else got, Kref) < 1e-13


def test_static_scratch_entry_point_rejects_sizes_beyond_its_limit(cuda):
    from botorch_b200 import _lib
    from botorch_b200.synthetic import make_problem

    p = make_problem("C2")
    st = gp_state(p, oracle_gp(p), cuda)
    Nc = STATIC_SCRATCH_DOUBLES // _lib.kp(p.d) + 8  # one padded column block too many
    Xc = torch.rand(Nc, p.d, dtype=torch.float64, device=cuda)
    code, _ = _static(st, Xc)
    assert code == 5  # BO_ERR_STATIC_SCRATCH
    with pytest.raises(_lib.BoError, match="built-in scratch"):
        _lib.check(code)
    # the largest size that fits still works
    Nc = STATIC_SCRATCH_DOUBLES // _lib.kp(p.d) // 8 * 8
    Xc = torch.rand(Nc, p.d, dtype=torch.float64, device=cuda)
    code, got = _static(st, Xc)
    _lib.check(code)
    assert torch.equal(got, st.kernel_matrix(Xc))
