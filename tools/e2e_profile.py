"""Host profile of the end-to-end evaluation (`optim.HostEvaluator`: the objective SciPy's L-BFGS-B calls).

    python tools/e2e_profile.py [--config C3] [--steps 100]

Prints the device-resident step time, the end-to-end time per evaluation, and a cProfile of the end-to-end loop sorted by
own time -- where the difference goes on the host."""
import argparse
import cProfile
import io
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C3")
    ap.add_argument("--steps", type=int, default=100)
    a = ap.parse_args()
    import bench
    import botorch_b200 as bb
    from botorch_b200.optim import HostEvaluator

    dev = torch.device("cuda:0")
    p = bench._problem(a.config)
    h = p.hypers
    if p.kernel == "rbf":
        model = bb.SingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev), lengthscale=h["lengthscale"], noise=h["noise"],
                                mean_constant=h["mean"], outcome_transform="DEFAULT" if p.standardize else None)  # fmt: skip
    else:
        model = bb.SaasFullyBayesianSingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev))
        model.load_mcmc_samples(h)
    acqf = bb.qLogNoisyExpectedImprovement(model, p.train_X.to(dev),
                                           sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234))  # fmt: skip
    X_host = bench._restart_batch(p, p.num_restarts)
    x_np = X_host.numpy().reshape(-1).copy()
    ev = HostEvaluator(acqf, X_host.shape, torch.empty(0, dtype=torch.float64, device=dev))
    for _ in range(5):
        ev(x_np)
    # device-resident: same Python path, no host copies, no per-step wait
    X = X_host.to(dev).requires_grad_(True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        loss = (-acqf(X)).sum()
        torch.autograd.grad(loss, X)
    torch.cuda.synchronize()
    t_dev = (time.perf_counter() - t0) / a.steps
    t0 = time.perf_counter()
    for _ in range(a.steps):
        ev(x_np)
    t_e2e = (time.perf_counter() - t0) / a.steps
    print(f"{a.config}: device-resident {1e3 * t_dev:.3f} ms/eval, end to end {1e3 * t_e2e:.3f} ms/eval")
    # phase timeline of one evaluation (host clock), averaged; GPU span from the H2D copy to the D2H copy by CUDA events
    import numpy as np

    names = ["x -> pinned", "H2D enqueue", "forward enqueue", "backward enqueue", "D2H enqueue", "host wait", "flags + copy out"]
    acc = np.zeros(len(names))
    gpu_span = 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(a.steps):
        t = [time.perf_counter()]
        ev.x_host.copy_(torch.from_numpy(x_np)); t.append(time.perf_counter())
        e0.record()
        Xd = ev.x_host.to(dev, non_blocking=True).view(ev.shapeX); t.append(time.perf_counter())
        Xd.requires_grad_(True)
        loss = (-acqf(Xd)).sum(); t.append(time.perf_counter())
        (grad,) = torch.autograd.grad(loss, Xd); t.append(time.perf_counter())
        ev.out_host.copy_(torch.cat([grad.reshape(-1), loss.detach().reshape(1)]), non_blocking=True)
        e1.record(); t.append(time.perf_counter())
        torch.cuda.current_stream(dev).synchronize(); t.append(time.perf_counter())
        acqf.check_numerics(quiet=True)
        out = ev.out_host.numpy()
        _ = float(out[-1]), out[:-1].astype(np.float64, copy=True); t.append(time.perf_counter())
        acc += np.diff(np.array(t))
        gpu_span += e0.elapsed_time(e1)
    for n_, v in zip(names, acc / a.steps):
        print(f"  {n_:18s} {1e3 * v:7.3f} ms")
    print(f"  GPU span H2D..D2H  {gpu_span / a.steps:7.3f} ms   (sum of host phases {1e3 * acc.sum() / a.steps:.3f} ms)")
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(a.steps):
        ev(x_np)
    pr.disable()
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(22)
    print(s.getvalue())


if __name__ == "__main__":
    main()
