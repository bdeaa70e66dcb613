"""Time bo_potrf (blocked Cholesky, optionally + triangular inverse) with CUDA events: single 64-block (the diagonal
kernel plus three tiny launches), n = 2000 / 8000, and the SAAS-shaped batch of 256 x (100 x 100)."""
import json
import sys

import torch

sys.path.insert(0, ".")
from botorch_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
lib, ctx = _lib.lib(), _lib.ctx(dev)


def run(n, batch, with_inv, reps):
    g = torch.Generator(device="cpu").manual_seed(n)
    B = torch.randn(batch, n, n, generator=g, dtype=torch.float64).to(dev)
    A0 = B @ B.mT / n + torch.eye(n, dtype=torch.float64, device=dev)
    ld = _lib.pad8(n)
    A = torch.zeros(batch, n, ld, dtype=torch.float64, device=dev)
    inv = torch.zeros(batch, n, ld, dtype=torch.float64, device=dev) if with_inv else None
    work = _lib.workspace(lib.bo_potrf_workspace_bytes(n, batch), dev)
    info = torch.zeros(batch, dtype=torch.int32, device=dev)
    src = torch.zeros_like(A)
    src[:, :, :n] = A0

    def once():
        A.copy_(src)
        _lib.check(lib.bo_potrf(ctx, n, batch, A.data_ptr(), ld, n * ld, None if inv is None else inv.data_ptr(),
                                work.data_ptr(), work.numel(), info.data_ptr(), _lib.stream_ptr(dev)))

    def copy_only():
        A.copy_(src)

    def t(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(reps):
            fn()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / reps

    ms = t(once) - t(copy_only)
    import time
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    host_ms = 1e3 * (time.perf_counter() - t0) / reps  # enqueue time only (no synchronisation)
    torch.cuda.synchronize()
    # the same call sequence replayed as a CUDA graph: device time without the host's launch cost
    gr = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        once()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    with torch.cuda.graph(gr):
        once()
    graph_ms = t(gr.replay) - t(copy_only)
    once()
    torch.cuda.synchronize()
    L = torch.linalg.cholesky(A0)
    err = float((A[:, :, :n] - L).abs().max() / L.abs().max())
    # the library beside it: torch.linalg.cholesky_ex (cuSOLVER potrf / batched) and, with the inverse, solve_triangular
    # against the identity (cuBLAS trsm), same matrices, same CUDA-event timing
    eye = torch.eye(n, dtype=torch.float64, device=dev).expand(batch, n, n).contiguous()

    def lib_once():
        Lc, _ = torch.linalg.cholesky_ex(A0)
        if with_inv:
            torch.linalg.solve_triangular(Lc, eye, upper=False)

    cusolver_ms = t(lib_once)
    flops = batch * (n**3 / 3.0 + (n**3 / 3.0 if with_inv else 0.0))
    r = {"n": n, "batch": batch, "inverse": with_inv, "ms": ms, "host_enqueue_ms": host_ms, "graph_replay_ms": graph_ms,
         "torch_cusolver_ms": cusolver_ms, "speedup_vs_torch": cusolver_ms / ms, "tflops_fp64": flops / (ms * 1e-3) / 1e12,
         "max_rel_err_vs_torch": err, "info": int(info.abs().max())}
    if with_inv:
        r["inv_residual"] = float((inv[:, :, :n] @ L - torch.eye(n, dtype=torch.float64, device=dev)).abs().max())
    return r


import os
if os.environ.get('BO_POTRF_MODE'):
    out = [run(2000, 1, False, 20)]
else:
    out = [run(64, 1, True, 200), run(2000, 1, False, 20), run(2000, 1, True, 20), run(100, 256, True, 50), run(8000, 1, True, 3),
           run(20000, 1, True, 2)]
print(json.dumps(out))
