"""Config C5: exact GP n=20000, d=12, marginal Thompson posterior over a candidate stream (single GPU share).
Usage: python tools/bench_c5.py [n] [candidates] [chunk]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import botorch_b200 as bb
from botorch_b200.synthetic import make_problem

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
Nc = int(sys.argv[2]) if len(sys.argv) > 2 else 131072
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 16384
dev = torch.device("cuda:0")
p = make_problem("C5", n=n)
h = p.hypers
t0 = time.perf_counter()
model = bb.SingleTaskGP(p.train_X.to(dev), p.train_Y.to(dev), lengthscale=h["lengthscale"], noise=h["noise"], mean_constant=h["mean"])
st = model.gp_state
torch.cuda.synchronize()
setup_s = time.perf_counter() - t0
ts = bb.MaxPosteriorSampling(model, chunk_size=chunk)
Xc = torch.quasirandom.SobolEngine(p.d, scramble=True, seed=1).draw(Nc, dtype=torch.float64).to(dev)
ts.marginal_posterior(Xc[:chunk])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
mean, var = ts.marginal_posterior(Xc)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
flops = Nc * (2 * (p.d + 2) * n + 2 * n + n * n + 2 * n)
g = torch.Generator(device=dev).manual_seed(0)
pick = ts(Xc[:chunk], num_samples=1, generator=g)
# spot parity on 64 candidates against torch fp64 on the GPU factors (oracle is too slow at n=20000)
L = st.L[0]; alpha = st.alpha[0]
xs = Xc[:64]
ls = h["lengthscale"].to(dev)
Kx = torch.exp(-0.5 * torch.cdist(p.train_X.to(dev) / ls, xs / ls) ** 2)
m_ref = float(st.y_mean) + st.y_std * (Kx.T @ alpha)
v = torch.linalg.solve_triangular(L, Kx, upper=False)
var_ref = st.y_std**2 * (1.0 - (v * v).sum(0))
print(json.dumps({"n": n, "candidates": Nc, "chunk": chunk, "setup_s": setup_s, "ms": ms, "cand_per_s": Nc / ms * 1e3,
                  "tflops_fp64": flops / ms * 1e-9, "mean_rel_err": float(((mean[0, :64] - m_ref).abs() / m_ref.abs().clamp_min(1e-3)).max()),
                  "var_err_vs_prior": float((var[0, :64] - var_ref).abs().max() / st.y_std**2), "jitter": st.jitter.tolist(),
                  "pick_shape": list(pick.shape)}))

# ---- pathwise (Matheron) Thompson sampling over the same candidate stream: joint draws, 16 paths ----------------------
torch.manual_seed(0)
t0 = time.perf_counter()
paths = bb.draw_matheron_paths(model, torch.Size([16]), num_features=1024)
paths.chunk_size = chunk
torch.cuda.synchronize()
path_setup_s = time.perf_counter() - t0
paths(Xc[:chunk])
torch.cuda.synchronize()
e0.record()
vals = paths(Xc)
best = vals.argmax(dim=-1)
e1.record(); torch.cuda.synchronize()
pms = e0.elapsed_time(e1)
# spot check against torch fp64 on the device factors: path = m + phi w + k(x, X) v
psx = paths.state
z = (xs / ls) @ psx.W.T
phi = (2.0 / psx.F) ** 0.5 * torch.cat([z.sin(), z.cos()], dim=-1)
ref = float(st.y_mean) + st.y_std * (float(h["mean"]) + psx.w @ phi.T + psx.v[:, :n] @ Kx)
print(json.dumps({"pathwise": {"paths": 16, "features": 1024, "candidates": Nc, "setup_s": path_setup_s, "ms": pms,
                               "cand_per_s": Nc / pms * 1e3, "path_values_per_s": 16 * Nc / pms * 1e3,
                               "rel_err_vs_torch_on_device_factors": float((vals[:, :64] - ref).abs().max() / ref.abs().max()),
                               "picked": best[:4].tolist()}}))
