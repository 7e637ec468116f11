"""Where does the host-buffer (e2e) loop lose time against the device-resident loop?  Run on a GPU box."""
import importlib
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "cs566-project-lightglue_b200"
M = importlib.import_module(PKG + ".matcher")
W = importlib.import_module(PKG + ".weights")
P = importlib.import_module(PKG + ".pipeline")

conf = {"filter_threshold": 0.1}
model = M.LightGlue(conf)
model.load_state_dict(W.make_state_dict(conf, seed=0))
model = model.cuda()
host = W.make_inputs(32, 2048, 2048, seed=1)
host = {k: ({kk: vv.pin_memory() for kk, vv in v.items()} if isinstance(v, dict) else v.pin_memory()) for k, v in host.items()}
data = {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in host.items()}
for _ in range(4):
    model(data)
torch.cuda.synchronize()


def gpu_ms(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


t0 = time.perf_counter()
for _ in range(10):
    model(data)
host_ms = (time.perf_counter() - t0) * 100
torch.cuda.synchronize()
print(f"host enqueue time per forward: {host_ms:.2f} ms")
print(f"resident: {gpu_ms(lambda: [model(data) for _ in range(20)]) / 20:.3f} ms/step")
for depth in (2, 3):
    pipe = P.HostPipeline(model, depth=depth)
    for steps in (10, 40):
        run = lambda: [None for _ in pipe.run(host for _ in range(steps))]
        run()
        print(f"pipeline depth {depth}, {steps} steps: {gpu_ms(run) / steps:.3f} ms/step")
# H2D alone
s = torch.cuda.Stream()
dst = torch.empty_like(data["descriptors0"])
print(f"H2D of one descriptor tensor: {gpu_ms(lambda: dst.copy_(host['descriptors0'], non_blocking=True)):.3f} ms")
