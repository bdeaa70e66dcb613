"""Sweep the DMMA GEMM tile configurations (BO_GEMM_CFG) on the C3 shapes, with torch.matmul (cuBLAS) timed on the same
shapes and every result checked against it.  Usage: python tools/gemm_bench.py [cfg ...]   (cfg -1 = the library's own
dispatch, 0-12 = cp.async variants, 20-24 = TMA-staged variants, "torch" = cuBLAS)
BO_SHAPES="M,N,K,tri;..." overrides the shape list (tri: 0 none, 1 A lower, 2 A upper, 3 B lower)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
KSWEEP = [("K%d" % k, 2000, 4096, k, 0) for k in (64, 128, 256, 512, 1024, 2000)]
SHAPES = [("dense 8192^3", 8192, 8192, 8192, 0), ("C3 fwd stacked (upper)", 2264, 4096, 2000, 2),
          ("C3 bwd stacked (lower)", 2000, 4096, 2264, 1), ("C3 dense 2264x4096x2000", 2264, 4096, 2000, 0),
          ("strong N=8 share 2264x512x2000 (upper)", 2264, 512, 2000, 2)]


def child():
    import torch
    from botorch_b200 import _lib

    dev = torch.device("cuda:0")
    lib = _lib.lib()
    ctx = _lib.ctx(dev)
    use_torch = os.environ.get("BO_TORCH") == "1"
    out = {}
    shapes = KSWEEP if os.environ.get("BO_KSWEEP") else SHAPES
    if os.environ.get("BO_SHAPES"):
        shapes = [("%sx%sx%s t%s" % tuple(t.split(",")), *map(int, t.split(","))) for t in os.environ["BO_SHAPES"].split(";")]
    for name, M, N, K, tri in shapes:
        lda, ldb = _lib.pad8(K), _lib.pad8(N)
        g = torch.Generator(device=dev).manual_seed(M + N)
        A = torch.randn(M, lda, dtype=torch.float64, device=dev, generator=g)
        B = torch.randn(K, ldb, dtype=torch.float64, device=dev, generator=g)
        if tri == 1:
            A = torch.tril(A)
        elif tri == 2:
            A = torch.triu(A)
        elif tri == 3:
            B = torch.tril(B)
        Cm = torch.empty(M, ldb, dtype=torch.float64, device=dev)

        def run():
            if use_torch:
                torch.matmul(A[:, :K], B[:, :N], out=Cm[:, :N]) if ldb == N else Cm[:, :N].copy_(A[:, :K] @ B[:, :N])
            else:
                _lib.check(lib.bo_dgemm(ctx, M, N, K, 1, 1.0, A.data_ptr(), lda, M * lda, B.data_ptr(), ldb, K * ldb, 0.0,
                                        Cm.data_ptr(), ldb, M * ldb, tri, _lib.stream_ptr(dev)))

        for _ in range(3):
            run()
        torch.cuda.synchronize()
        ref = A[:, :K] @ B[:, :N]
        err = float((Cm[:, :N] - ref).abs().max() / ref.abs().max())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5 if M * N * K > 1e11 else 20
        e0.record()
        for _ in range(reps):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        fl = 2.0 * M * N * K * (0.5 if tri else 1.0)  # nominal: triangular operands count half
        out[name] = {"ms": round(ms, 4), "tflops_nominal": round(fl / ms * 1e-9, 2), "rel_err_vs_torch": err}
    print(json.dumps(out))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        child()
    else:
        cfgs = sys.argv[1:] or ["torch", "-1", "10", "20", "21", "22", "23", "24"]
        for c in cfgs:
            env = dict(os.environ)
            if c == "torch":
                env["BO_TORCH"] = "1"
            else:
                env["BO_GEMM_CFG"] = c
            r = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True)
            print("cfg", c, r.stdout.strip() or r.stderr[-600:], flush=True)
