"""Runs, once each and at representative sizes, the kernels that are NOT part of the qLogNEI step, inside a
cudaProfilerStart/Stop window, so that `ncu --set full --profile-from-start off --kernel-id :::1` records one launch of each:

  * posterior API: joint posterior + its reverse (n = 2000, 64 x q = 8), marginal mean/variance of 16 384 candidates
    (the C5 scoring chunk, here against n = 4000 so that the capture stays short), Thompson argmax;
  * pathwise Thompson sampling: Matheron path construction and evaluation (1024 features);
  * the general-baseline kernel gradient (`bo_kernel_matrix_bwd`), constrained reductions (`bo_mc_reduce_*`);
  * the device L-BFGS kernels (C2, 5 rounds); the MLL value + gradient (n = 2000); the SAAS potential (64 chains).

    ncu --set full --profile-from-start off --kernel-id :::1 -o /tmp/other python tools/other_kernels_once.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    import bench
    import botorch_b200 as bb
    from botorch_b200.optim import gen_candidates_lbfgs
    from botorch_b200.synthetic import make_problem

    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream(dev)
    p3 = make_problem("C3")
    h = p3.hypers
    kw = dict(lengthscale=h["lengthscale"], noise=h["noise"], mean_constant=h["mean"])
    model = bb.SingleTaskGP(p3.train_X.to(dev), p3.train_Y.to(dev), **kw)
    g = torch.Generator().manual_seed(0)
    Xq = torch.rand(64, 8, p3.d, generator=g, dtype=torch.float64).to(dev)
    p5 = make_problem("C5", n=4000)
    h5 = p5.hypers
    model5 = bb.SingleTaskGP(p5.train_X.to(dev), p5.train_Y.to(dev), lengthscale=h5["lengthscale"], noise=h5["noise"],
                             mean_constant=h5["mean"])  # fmt: skip
    Xc = torch.rand(16384, p5.d, generator=g, dtype=torch.float64).to(dev)
    ts = bb.MaxPosteriorSampling(model5, chunk_size=16384)
    ts.marginal_posterior(Xc[:256])  # factor the training set outside the window
    p2 = make_problem("C2")
    h2 = p2.hypers
    model2 = bb.SingleTaskGP(p2.train_X.to(dev), p2.train_Y.to(dev), lengthscale=h2["lengthscale"], noise=h2["noise"],
                             mean_constant=h2["mean"])  # fmt: skip
    acq2 = bb.qLogNoisyExpectedImprovement(model2, p2.train_X.to(dev),
                                           sampler=bb.SobolQMCNormalSampler(torch.Size([p2.mc_samples]), seed=1))  # fmt: skip
    ics = torch.rand(p2.num_restarts, p2.q, p2.d, generator=g, dtype=torch.float64).to(dev)
    torch.cuda.synchronize()

    torch.cuda.profiler.start()
    # posterior API (A8/A16) and its reverse
    X = Xq.clone().requires_grad_(True)
    post = model.posterior(X)
    (post.mean.sum() + post.mvn.covariance_matrix.sum()).backward()
    # C5 scoring chunk + Thompson argmax (T1)
    ts(Xc, num_samples=2)
    # pathwise TS (N4)
    pts = bb.PathwiseThompsonSampling(model, num_features=1024)
    Xp = Xq.clone().requires_grad_(True)
    pts(Xp).sum().backward()
    # device L-BFGS (N3)
    gen_candidates_lbfgs(ics, acq2, lower_bounds=0.0, upper_bounds=1.0, options={"maxiter": 5})
    # MLL value + gradient (N4), constrained reductions (N2), SAAS potential (N4)
    bench._mll_leg(p3, dev, stream, cpu=False)
    bench._constrained_leg(p3, dev, 64)
    bench._saas_leg(dev)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("done")


if __name__ == "__main__":
    main()
