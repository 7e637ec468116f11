"""Experiment: accuracy / speed of the forward when the cuBLAS linears run TF32 vs fp32."""
import importlib, sys, os, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import lightglue_oracle as O
PKG = "cs566-project-lightglue_b200"
W = importlib.import_module(PKG + ".weights"); M = importlib.import_module(PKG + ".matcher")
conf = {"filter_threshold": 0.1}
sd = W.make_state_dict(conf, seed=0)
data = W.make_inputs(1, 2048, 2048, seed=1, correlated=0.7)
ref = O.forward(sd, conf, data)
model = M.LightGlue(conf); model.load_state_dict(sd); model = model.cuda()
cu = lambda d: {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in d.items()}
big = cu(W.make_inputs(32, 2048, 2048, seed=1))
for tf32 in (False, True):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    p = model(cu(data))
    err = (p["log_assignment"].cpu() - ref["log_assignment"]).abs().max().item()
    agree = (p["matches0"].cpu() == ref["matches0"]).float().mean().item()
    for _ in range(2): model(big)
    torch.cuda.synchronize(); t = time.time()
    for _ in range(3): model(big)
    torch.cuda.synchronize(); dt = (time.time() - t) / 3
    print(f"tf32={tf32}: err={err:.3e} agree={agree:.4f} step={dt*1e3:.1f} ms -> {32/dt:.0f} pairs/s")
