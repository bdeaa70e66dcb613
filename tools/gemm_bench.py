"""Sweep the DMMA GEMM tile configurations (BO_GEMM_CFG) on the C3 shapes. Usage: python tools/gemm_bench.py"""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
KSWEEP = [("K%d" % k, 2000, 4096, k, 0) for k in (64, 128, 256, 512, 1024, 2000)]
SHAPES = [("V tri", 2001, 4096, 2000, 1), ("Sigma", 2000, 4096, 2000, 0), ("F", 256, 4096, 2000, 0), ("Kbar upper", 2000, 4096, 2001, 2)]

def child():
    import torch
    from botorch_b200 import _lib
    dev = torch.device("cuda:0"); lib = _lib.lib(); ctx = _lib.ctx(dev)
    out = {}
    shapes = KSWEEP if os.environ.get('BO_KSWEEP') else SHAPES
    if os.environ.get('BO_SHAPES'):  # "M,N,K,tri;M,N,K,tri"
        shapes = [("%sx%sx%s t%s" % tuple(t.split(",")), *map(int, t.split(","))) for t in os.environ['BO_SHAPES'].split(";")]
    for name, M, N, K, tri in shapes:
        lda, ldb = _lib.pad8(K), _lib.pad8(N)
        A = torch.randn(M, lda, dtype=torch.float64, device=dev); B = torch.randn(K, ldb, dtype=torch.float64, device=dev)
        C = torch.empty(M, ldb, dtype=torch.float64, device=dev)
        def run():
            _lib.check(lib.bo_dgemm(ctx, M, N, K, 1, 1.0, A.data_ptr(), lda, M * lda, B.data_ptr(), ldb, K * ldb, 0.0,
                                    C.data_ptr(), ldb, M * ldb, tri, _lib.stream_ptr(dev)))
        for _ in range(3): run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): run()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        fl = 2.0 * M * N * K * (0.5 if tri else 1.0)
        out[name] = (round(ms, 4), round(fl / ms * 1e-9, 2))
    print(json.dumps(out))

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        child()
    else:
        cfgs = sys.argv[1:] or [str(i) for i in range(-1, 8)]
        for c in cfgs:
            env = dict(os.environ, BO_GEMM_CFG=c)
            r = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True)
            print("cfg", c, r.stdout.strip() or r.stderr[-400:], flush=True)
