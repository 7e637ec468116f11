"""Experiment (CPU, oracle): log_assignment error when the per-token linears use reduced-precision
tensor-core operand schemes.  Decides the operand format of lg_linear."""
import importlib, sys, os
import torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import lightglue_oracle as O
W = importlib.import_module("cs566-project-lightglue_b200.weights")

def rnd(x, bits):  # round-to-nearest-even to `bits` explicit mantissa bits
    u = x.contiguous().view(torch.int32)
    drop = 23 - bits
    u = (u + ((1 << (drop - 1)) - 1) + ((u >> drop) & 1)) & ~((1 << drop) - 1)
    return u.view(torch.float32)

def make_lin(mode):
    def lin(sd, name, x):
        w, b = sd[name + ".weight"], sd.get(name + ".bias")
        if mode == "fp32" or w.shape[0] == 1 or w.shape[1] < 16:
            return F.linear(x, w, b)
        if mode == "tf32":
            return F.linear(rnd(x, 10), rnd(w, 10), b)
        if mode == "bf16":
            return F.linear(rnd(x, 7), rnd(w, 7), b)
        if mode == "bf16x3":
            xh, wh = rnd(x, 7), rnd(w, 7)
            xl, wl = rnd(x - xh, 7), rnd(w - wh, 7)
            return F.linear(xh, wh, b) + (F.linear(xl, wh) + F.linear(xh, wl))
        if mode == "fp16x2w":  # activation split, weights single tf32-grade
            xh, wh = rnd(x, 10), rnd(w, 10)
            xl = rnd(x - xh, 10)
            return F.linear(xh, wh, b) + F.linear(xl, wh)
        raise ValueError(mode)
    return lin

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
conf = {"filter_threshold": 0.1}
sd = W.make_state_dict(conf, seed=0)
data = W.make_inputs(1, n, n, seed=1, correlated=0.7)
ref = O.forward(sd, conf, data)
orig = O._lin
for mode in ("tf32", "bf16x3", "fp16x2w", "bf16"):
    O._lin = make_lin(mode)
    p = O.forward(sd, conf, data)
    O._lin = orig
    err = (p["log_assignment"] - ref["log_assignment"]).abs().max().item()
    agree = (p["matches0"] == ref["matches0"]).float().mean().item()
    print(f"N={n} linears={mode:8s} max|dlog_assignment|={err:.3e} matches0 agreement={agree:.4f}", flush=True)
