"""fp32 path (b200.precision=fp32): one warm forward + one forward of B pairs x N kpts, for an ncu launch list:
   ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/fp32_launches.csv \
       python scripts/profile_fp32_path.py 4 2048"""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "cs566-project-lightglue_b200"
M = importlib.import_module(PKG + ".matcher")
W = importlib.import_module(PKG + ".weights")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
conf = {"filter_threshold": 0.1, "b200": {"precision": "fp32"}}
model = M.LightGlue(conf)
model.load_state_dict(W.make_state_dict(conf, seed=0))
model = model.cuda()
data = {k: (v.cuda() if torch.is_tensor(v) else {kk: vv.cuda() for kk, vv in v.items()}) for k, v in
        W.make_inputs(B, N, N, seed=1, correlated=0.7).items()}
for _ in range(2):
    model(data)
torch.cuda.synchronize()
