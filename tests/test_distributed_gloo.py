"""world_size = 2 on the gloo backend (CPU): the N > 1 host logic of botorch_b200.distributed -- block sharding of raw
samples with global index order, all-gather, identical initial conditions on every rank, best-candidate gather with
first-occurrence tie-break, factor broadcast plumbing."""
import os
import socket
import tempfile

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _Quad(torch.nn.Module):
    """Closed-form stand-in for an acquisition function (host tensors)."""

    def forward(self, X):
        t = torch.tensor([0.25, 0.6, 0.9], dtype=X.dtype)
        return -((X - t) ** 2).sum(dim=(-1, -2))


def _worker(rank: int, ws: int, port: int, outdir: str) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        from botorch_b200 import distributed as bdist
        from botorch_b200.sampling import draw_sobol_samples

        assert bdist.world() == (rank, ws)
        bounds = torch.stack([torch.zeros(3), torch.ones(3)]).double()
        acq = _Quad()
        # raw samples: each rank evaluates its block; the gathered vector equals the single-process evaluation
        X_rnd = draw_sobol_samples(bounds, n=37, q=2, seed=3)  # 37: uneven blocks (19 + 18)
        y = bdist.sharded_raw_sample_values(acq, X_rnd, torch.device("cpu"), batch_limit=5)
        assert torch.equal(y, acq(X_rnd))
        # broadcast of "factors"
        buf = [torch.full((4, 4), float(rank + 1), dtype=torch.float64), torch.arange(3, dtype=torch.float64) * (rank + 1)]
        nbytes = bdist.broadcast_buffers(buf, src=0)
        assert nbytes == 16 * 8 + 3 * 8 and float(buf[0][0, 0]) == 1.0 and float(buf[1][2]) == 2.0
        # best gather: equal maxima on both ranks -> the lower global index wins
        vals = torch.tensor([0.0, 3.0, 1.0], dtype=torch.float64) if rank == 0 else torch.tensor([3.0, 2.0], dtype=torch.float64)
        cands = torch.full((len(vals), 2, 3), float(rank), dtype=torch.float64)
        lo = 0 if rank == 0 else 3
        cand, val, idx = bdist.gather_best(vals, cands, lo)
        assert idx == 1 and float(val) == 3.0 and float(cand[0, 0]) == 0.0
        # sharded optimize_acqf: both ranks return the same winner, close to the optimum
        cand, val = bdist.optimize_acqf_sharded(acq, bounds, q=2, num_restarts=6, raw_samples=32,
                                                options={"seed": 0, "maxiter": 100})  # fmt: skip
        # each rank draws only ITS rows of the raw samples (the engine fast-forwarded to its block) and the selected rows are
        # fetched by an all-gather: with an "optimiser" that returns its starting points, the winner must be exactly the
        # best of the initial conditions a single process selects from the full draw (uneven blocks: 37 = 19 + 18)
        def identity(ics, acqf, lower_bounds=None, upper_bounds=None, options=None):
            return ics, acqf(ics)

        cand_i, val_i = bdist.optimize_acqf_sharded(acq, bounds, q=2, num_restarts=5, raw_samples=37, options={"seed": 3},
                                                    gen_candidates=identity)  # fmt: skip
        # no seed given: rank 0 draws it and broadcasts it, although the ranks' global RNGs differ (every rank used
        # to scramble its own Sobol sequence, so the gathered values belonged to other ranks' points)
        torch.manual_seed(100 + rank)
        seeds = [bdist.shared_seed(None) for _ in range(2)]
        assert bdist.shared_seed(17) == 17
        timings = {}
        cand_u, val_u = bdist.optimize_acqf_sharded(acq, bounds, q=2, num_restarts=6, raw_samples=32,
                                                    options={"maxiter": 100}, timings=timings)  # fmt: skip
        assert set(timings) == {"raw_sample_draw_s", "raw_sample_eval_and_allgather_s", "initial_condition_selection_s",
                                "restart_optimisation_s", "best_gather_s"}  # fmt: skip
        bdist.warm_up()
        # mixed dtypes travel as one packed buffer per dtype
        mixed = [torch.full((5,), float(rank), dtype=torch.float64), torch.full((3,), rank + 7, dtype=torch.int32),
                 torch.full((2, 2), float(rank) - 1.0, dtype=torch.float64)]  # fmt: skip
        bdist.broadcast_buffers(mixed, src=0)
        assert float(mixed[0][4]) == 0.0 and int(mixed[1][0]) == 7 and float(mixed[2][1, 1]) == -1.0
        # NUTS chains sharded over the ranks: distinct chains per rank, one all-gather, same ensemble everywhere
        mu = torch.tensor([1.0, -1.0], dtype=torch.float64)

        def make_potential(num_chains):
            return lambda z: (0.5 * ((z - mu) ** 2).sum(-1), z - mu)

        draws, _ = bdist.nuts_sample_sharded(make_potential, 2, 3, num_samples=40, warmup_steps=60, device=torch.device("cpu"),
                                             seed=5)  # fmt: skip
        assert draws.shape == (40, 6, 2)
        torch.save({"cand": cand, "val": val, "draws": draws, "cand_u": cand_u, "val_u": val_u, "seeds": seeds,
                    "cand_i": cand_i, "val_i": val_i}, os.path.join(outdir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    port = _free_port()
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(2, port, d), nprocs=2, join=True)
        a, b = torch.load(os.path.join(d, "r0.pt")), torch.load(os.path.join(d, "r1.pt"))
    assert torch.equal(a["cand"], b["cand"]) and torch.equal(a["val"], b["val"])
    assert torch.allclose(a["cand"], torch.tensor([0.25, 0.6, 0.9], dtype=torch.float64).expand(2, 3), atol=1e-4)
    assert a["seeds"] == b["seeds"] and a["seeds"][0] != a["seeds"][1]
    assert torch.equal(a["cand_u"], b["cand_u"]) and torch.equal(a["val_u"], b["val_u"])
    assert torch.allclose(a["cand_u"], torch.tensor([0.25, 0.6, 0.9], dtype=torch.float64).expand(2, 3), atol=1e-4)
    # row-sharded raw samples: the single-process selection from the full draw, same seed, same host RNG protocol
    from botorch_b200.optim import initialize_q_batch_indices
    from botorch_b200.sampling import draw_sobol_samples

    bounds = torch.stack([torch.zeros(3), torch.ones(3)]).double()
    X_full = draw_sobol_samples(bounds, n=37, q=2, seed=3)
    Y_full = _Quad()(X_full)
    state = torch.random.get_rng_state()
    torch.manual_seed(3)
    ics = X_full[initialize_q_batch_indices(Y_full, 5)]
    torch.random.set_rng_state(state)
    best = int(torch.argmax(_Quad()(ics)))
    assert torch.equal(a["cand_i"], ics[best]) and torch.equal(b["cand_i"], ics[best])
    assert float(a["val_i"]) == float(_Quad()(ics)[best])
    assert torch.equal(a["draws"], b["draws"])  # every rank holds the same gathered chains ...
    chains = a["draws"].transpose(0, 1)
    assert all(not torch.equal(chains[i], chains[j]) for i in range(6) for j in range(i))  # ... and the chains are distinct
    assert float((a["draws"].reshape(-1, 2).mean(0) - torch.tensor([1.0, -1.0])).abs().max()) < 0.3
