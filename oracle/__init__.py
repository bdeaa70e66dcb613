"""CPU oracle for the qLogNEI / GP-posterior hot path of anand-12/botorch.

TEST INFRASTRUCTURE ONLY.  Nothing under ``botorch_b200/`` imports this package.
The only legal importers are ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``; there it is the
checker or the timed CPU baseline, never the product path.

What it is: a float64, pure-torch (CPU) restatement of the reference's algorithm
for this path, written function by function against the reference file:line
each docstring cites.  The reference is pure Python; its linear algebra lives
in two third-party packages that are NOT vendored under /root/reference and are
not installable here: gpytorch==1.12 and linear_operator==0.5.2
(requirements.txt:6-7).  Their published algorithms (kernel evaluation through
the mean-centred quadratic-expansion distance, Cholesky prediction caches,
the fast_pred_var ``L^-T`` GEMM covariance, ``psd_safe_cholesky`` jitter retry,
``MultivariateNormal.rsample``) are restated in ``oracle/gp.py``.

Pinning status
  * reductions (log_fatplus / fatmax / logmeanexp), Sobol-normal base samples,
    raw Sobol candidates, synthetic test functions and the t-batch decorators:
    PINNED.  ``tests/golden/make_golden.py`` imports the reference's own
    modules verbatim from /root/reference (botorch/utils/safe_math.py,
    botorch/sampling/qmc.py, botorch/utils/sampling.py,
    botorch/test_functions/synthetic.py) and writes the committed fixtures the
    oracle (and the CUDA path) are checked against, together with the
    known-answer cases of test/acquisition/test_logei.py and
    test/utils/test_safe_math.py.
  * GP posterior values (K1-K4 of SURVEY.md section 2c): PARITY UNPINNED by reference
    fixtures -- the reference's tests hold no golden vectors for them and
    gpytorch cannot be imported here.  They are pinned only by (a) two
    independent formulations inside the oracle (direct ``K - Kx K^-1 Kx^T``
    solve vs. the ``L^-T`` GEMM form), (b) the reference's own self-consistency
    identities re-created as tests (cached == uncached sampling, test_low_rank
    / test_logei.test_cache_root).
"""
