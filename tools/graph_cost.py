"""What a CUDA-graph capture of one fused forward+backward costs on the host, and what a replay saves per evaluation.

    python tools/graph_cost.py [--config C3] [--batches 4,16,32]

For each t-batch count B: wall time of `GraphedEval(...)` (three warm-up runs + capture + instantiate), wall time per
eager evaluation and per replay, each over 50 back-to-back evaluations with one synchronise at the end (the L-BFGS loop
does not synchronise per round).  Break-even uses = capture / (eager - replay).
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C3")
    ap.add_argument("--batches", default="4,16,32")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import bench
    from botorch_b200.engine import GraphedEval

    dev = torch.device("cuda:0")
    import botorch_b200 as bb

    p = bench._problem(a.config)
    h = p.hypers
    model = bb.SingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev), lengthscale=h["lengthscale"], noise=h["noise"],
                            mean_constant=h["mean"], outcome_transform="DEFAULT" if p.standardize else None)  # fmt: skip
    acqf = bb.qLogNoisyExpectedImprovement(model, p.train_X.to(dev),
                                           sampler=bb.SobolQMCNormalSampler(torch.Size([p.mc_samples]), seed=1234))  # fmt: skip
    acqf._set_sampler(p.q)
    nei, Zq, taus = acqf._nei, acqf._Zq, (acqf.tau_relu, acqf.tau_max)
    rows = []
    for B in [int(b) for b in a.batches.split(",")]:
        X = torch.rand(B, p.q, p.d, dtype=torch.float64, device=dev)
        w = torch.ones(nei.gp.M, B, dtype=torch.float64, device=dev)

        def eager():
            acq, info, work = nei.forward(X, Zq, taus[0], taus[1], with_grad=True)
            return nei.backward(X, Zq, taus[0], taus[1], w, work)

        for _ in range(3):
            eager()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(50):
            eager()
        torch.cuda.synchronize()
        t_eager = (time.perf_counter() - t0) / 50
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ge = GraphedEval(nei, B, p.q, p.d, Zq, taus[0], taus[1], True)
        torch.cuda.synchronize()
        t_cap = time.perf_counter() - t0
        for _ in range(3):
            ge.run_forward(X)
            ge.run_backward(w)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(50):
            ge.run_forward(X)
            ge.run_backward(w)
        torch.cuda.synchronize()
        t_replay = (time.perf_counter() - t0) / 50
        gain = t_eager - t_replay
        rows.append({"config": a.config, "B": B, "capture_ms": 1e3 * t_cap, "eager_ms": 1e3 * t_eager,
                     "replay_ms": 1e3 * t_replay, "break_even_uses": (t_cap / gain) if gain > 0 else None})  # fmt: skip
        print(json.dumps(rows[-1]), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
