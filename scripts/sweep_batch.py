"""BASELINE config 5 on one GPU: pairs per step 8..256 at 2048 kpts (device-resident inputs), plus 4096 / 8192 keypoints
sanity points.  One JSON line per point."""
import importlib, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "cs566-project-lightglue_b200"
W = importlib.import_module(PKG + ".weights"); M = importlib.import_module(PKG + ".matcher")
conf = {"filter_threshold": 0.1}
sd = W.make_state_dict(conf, seed=0)
model = M.LightGlue(conf); model.load_state_dict(sd); model = model.cuda()


def cuda(data):
    return {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in data.items()}


def point(B, N, steps=5):
    data = cuda(W.make_inputs(B, N, N, seed=1))
    for _ in range(3):
        pred = model(data)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        pred = model(data)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    ok = bool(torch.isfinite(pred["log_assignment"]).all())
    print(json.dumps({"pairs": B, "kpts": N, "ms_per_step": round(ms, 3), "pairs_per_s": round(B / ms * 1e3, 1), "finite": ok,
                      "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2)}), flush=True)
    del pred, data
    torch.cuda.empty_cache()


for B, N in ((1, 1024), (1, 2048), (2, 2048), (4, 2048)):  # the reference's exporters call the matcher one pair at a time
    point(B, N, steps=20)
if os.environ.get("LG_SWEEP_SMALL"):
    sys.exit(0)
for B in (8, 16, 32, 64, 128, 256):
    point(B, 2048)
point(8, 4096)
point(2, 8192)
