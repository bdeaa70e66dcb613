"""Two bo_mll_value_grad evaluations at a BASELINE shape (profiling target: ncu launch lists / full captures)."""
import sys

import torch

sys.path.insert(0, ".")
from botorch_b200 import engine  # noqa: E402
from botorch_b200.synthetic import make_problem  # noqa: E402

dev = torch.device("cuda:0")
p = make_problem(sys.argv[1] if len(sys.argv) > 1 else "C3")
y = ((p.train_Y - p.train_Y.mean()) / p.train_Y.std()).reshape(-1)
st = engine.MLLState(p.train_X.to(dev), y.to(dev), p.kernel, with_outputscale=False)
h = p.hypers
for _ in range(2):
    v, g = st.evaluate(h["lengthscale"].reshape(-1).to(dev), h["noise"].reshape(()).to(dev), h["mean"].reshape(()).to(dev))
torch.cuda.synchronize()
print(float(v[0]), g[0, :3].tolist())
