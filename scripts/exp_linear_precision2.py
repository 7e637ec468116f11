"""Experiment (CPU, oracle): which of the per-token linears need the 3-term split?  Each group of linears in turn is
run with fp16-grade operands (10 explicit mantissa bits, ONE product) while every other linear keeps the built
3-term split bf16; the end-to-end error of log_assignment against the fp32 oracle decides (precision table).
In the built engine out_proj / to_out are folded into ffn.0 (engine.py), so "ffn0" covers them too."""
import importlib, sys, os
import torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import lightglue_oracle as O
W = importlib.import_module("cs566-project-lightglue_b200.weights")


def rnd(x, bits):
    u = x.contiguous().view(torch.int32)
    drop = 23 - bits
    u = (u + ((1 << (drop - 1)) - 1) + ((u >> drop) & 1)) & ~((1 << drop) - 1)
    return u.view(torch.float32)


GROUPS = {
    "qkv": ("Wqkv", "to_qk", "to_v"),
    "ffn0": ("ffn.0", "out_proj", "to_out"),
    "ffn3": ("ffn.3",),
    "final": ("final_proj",),
    "input": ("input_proj",),
}


def make_lin(single):
    def lin(sd, name, x):
        w, b = sd[name + ".weight"], sd.get(name + ".bias")
        if w.shape[0] == 1 or w.shape[1] < 16:
            return F.linear(x, w, b)
        if any(name.endswith(s) for g in single for s in GROUPS[g]):
            return F.linear(rnd(x, 10), rnd(w, 10), b)
        xh, wh = rnd(x, 7), rnd(w, 7)
        xl, wl = rnd(x - xh, 7), rnd(w - wh, 7)
        return F.linear(xh, wh, b) + (F.linear(xl, wh) + F.linear(xh, wl))
    return lin


n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
torch.set_num_threads(os.cpu_count() or 8)
conf = {"filter_threshold": 0.1}
sd = W.make_state_dict(conf, seed=0)
data = W.make_inputs(1, n, n, seed=1, correlated=0.7)
ref = O.forward(sd, conf, data)
orig = O._lin
cases = [(), ("qkv",), ("ffn0",), ("ffn3",), ("final",), ("input",), ("qkv", "ffn0"), ("qkv", "ffn0", "ffn3"),
         ("qkv", "ffn0", "ffn3", "final", "input")]
for single in cases:
    O._lin = make_lin(single)
    p = O.forward(sd, conf, data)
    O._lin = orig
    err = (p["log_assignment"] - ref["log_assignment"]).abs().max().item()
    agree = (p["matches0"] == ref["matches0"]).float().mean().item()
    print(f"N={n} single-product fp16 in {'+'.join(single) or 'none':28s} max|dlog_assignment|={err:.3e} "
          f"matches0 agreement={agree:.4f}", flush=True)
