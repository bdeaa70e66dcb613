"""fit_fully_bayesian_model_nuts with the chains sharded over the ranks (one process per GPU, NCCL):
python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/saas_sharded.py"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, ".")
import botorch_b200 as bb  # noqa: E402
from botorch_b200 import distributed as bdist  # noqa: E402

rank, ws, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if ws > 1:
    dist.init_process_group("nccl", device_id=dev)
g = torch.Generator().manual_seed(0)
n, d = 100, 30
X = torch.rand(n, d, generator=g, dtype=torch.float64)
y = torch.sin(2 * torch.pi * X[:, 0]) + 0.1 * torch.randn(n, generator=g, dtype=torch.float64)
y = ((y - y.mean()) / y.std()).unsqueeze(-1)
model = bb.SaasFullyBayesianSingleTaskGP(X.to(dev), y.to(dev))
torch.cuda.synchronize()
if ws > 1:
    dist.barrier()
t0 = time.perf_counter()
bdist.fit_fully_bayesian_model_nuts_sharded(model, warmup_steps=512, num_samples=256 * ws, thinning=16, chains_per_rank=4, seed=0)
torch.cuda.synchronize()
dt = torch.tensor([time.perf_counter() - t0], device=dev)
check = torch.stack([model._mcmc["lengthscale"].sum(), model._mcmc["noise"].sum()])
if ws > 1:
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    allc = [torch.empty_like(check) for _ in range(ws)]
    dist.all_gather(allc, check)
    same = all(torch.equal(allc[0], c) for c in allc)
else:
    same = True
if rank == 0:
    print(json.dumps({"n_gpus": ws, "chains_total": 4 * ws, "retained_samples": model.num_mcmc_samples, "seconds_max_over_ranks": float(dt),
                      "same_ensemble_on_every_rank": same, "median_lengthscale_dim0": float(model.median_lengthscale[0]),
                      "median_lengthscale_others_min": float(model.median_lengthscale[1:].min()),
                      "accept_prob_rank0": float(model.nuts_stats["accept_prob"].mean())}))
if ws > 1:
    dist.destroy_process_group()
