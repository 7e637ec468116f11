"""Experiment: Wqkv / to_qk|to_v as single hi.hi bf16 products (LG_QKV_TERMS=1) vs the 3-term split: error of
log_assignment against the oracle / goldens at several sizes, match agreement, and the bench-shape step time."""
import importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_golden
from oracle import lightglue_oracle as O
W = importlib.import_module("cs566-project-lightglue_b200.weights")
M = importlib.import_module("cs566-project-lightglue_b200.matcher")
cuda = lambda d: {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in d.items()}
print("LG_QKV_TERMS =", os.environ.get("LG_QKV_TERMS", "3"))
for name in ["sp_b1_192", "sp_b2_160x224", "disk_b1_130x97", "sp_b1_1024"]:
    meta, g = load_golden("fwd_" + name)
    conf = meta["conf"]
    sd = W.make_state_dict(conf, seed=meta["weight_seed"])
    data = W.make_inputs(meta["batch"], meta["m"], meta["n"], d_in=conf.get("input_dim", 256), seed=meta["data_seed"],
                         image_size=tuple(meta["image_size"]), correlated=meta["correlated"])
    m = M.LightGlue(conf); m.load_state_dict(sd); m = m.cuda()
    pred = m(cuda(data))
    la = pred["log_assignment"].cpu().numpy()
    err = np.abs(la - g["log_assignment"]).max() if "log_assignment" in g else np.abs(la[:, ::17, ::19] - g["log_assignment_sample"]).max()
    print(f"{name:18s} err {err:.2e} agree {(pred['matches0'].cpu().numpy() == g['matches0']).mean():.4f}")
conf = {"filter_threshold": 0.0}
sd = W.make_state_dict(conf, seed=0)
m = M.LightGlue(conf); m.load_state_dict(sd); m = m.cuda()
for n in (2048, 4096):
    data = W.make_inputs(1, n, n, seed=1, correlated=0.7)
    ref = O.forward(sd, conf, data)
    pred = m(cuda(data))
    print(f"{n}: err {(pred['log_assignment'].cpu() - ref['log_assignment']).abs().max():.2e} "
          f"agree {(pred['matches0'].cpu() == ref['matches0']).float().mean():.4f}")
data = cuda(W.make_inputs(32, 2048, 2048, seed=1))
for _ in range(5): m(data)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20): m(data)
torch.cuda.synchronize(); print(f"32 x 2048: {(time.perf_counter() - t0) / 20 * 1e3:.2f} ms/step")
