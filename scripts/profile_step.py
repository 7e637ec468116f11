"""One profiled forward of the bench workload (32 pairs x 2048 kpts) between
cudaProfilerStart/Stop, for `ncu --profile-from-start off`."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "cs566-project-lightglue_b200"
W = importlib.import_module(PKG + ".weights"); M = importlib.import_module(PKG + ".matcher")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
conf = {"filter_threshold": 0.1}
model = M.LightGlue(conf); model.load_state_dict(W.make_state_dict(conf, seed=0)); model = model.cuda()
d = W.make_inputs(B, N, N, seed=1)
d = {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in d.items()}
for _ in range(2):
    model(d)
torch.cuda.synchronize()
torch.cuda.profiler.start()
model(d)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
