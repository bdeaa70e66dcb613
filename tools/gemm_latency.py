"""Per-launch time of small-K DMMA GEMMs (the panel / trailing shapes of the blocked Cholesky), forced tile configs."""
import json
import os
import subprocess
import sys

if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch

    sys.path.insert(0, ".")
    from botorch_b200 import _lib

    dev = torch.device("cuda:0")
    lib, ctx = _lib.lib(), _lib.ctx(dev)
    res = {}
    for (M, N, K) in [(1984, 64, 128), (1984, 64, 64), (1984, 1984, 64), (960, 960, 64), (64, 64, 64)]:
        lda, ldb, ldc = _lib.pad8(K), _lib.pad8(N), _lib.pad8(N)
        A = torch.randn(M, lda, dtype=torch.float64, device=dev)
        B = torch.randn(K, ldb, dtype=torch.float64, device=dev)
        Cm = torch.zeros(M, ldc, dtype=torch.float64, device=dev)

        def f():
            _lib.check(lib.bo_dgemm(ctx, M, N, K, 1, 1.0, A.data_ptr(), lda, M * lda, B.data_ptr(), ldb, K * ldb, 0.0,
                                    Cm.data_ptr(), ldc, M * ldc, 0, _lib.stream_ptr(dev)))

        for _ in range(5):
            f()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()  # 100 dependent launches in one graph: device time per launch without host cost
        with torch.cuda.graph(gr):
            for _ in range(100):
                f()
        gr.replay()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(5):
            gr.replay()
        e.record()
        torch.cuda.synchronize()
        res[f"{M}x{N}x{K}"] = round(s.elapsed_time(e) / 500 * 1e3, 2)
    print(json.dumps(res))
else:
    out = {}
    for cfg in ["-1", "10", "12"]:
        env = dict(os.environ, BO_GEMM_CFG=cfg)
        if cfg == "-1":
            env.pop("BO_GEMM_CFG")
        r = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True)
        out[cfg] = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else r.stderr[-300:]
    print(json.dumps(out, indent=1))
