"""One pair per call, adaptive depth + width at 2048 keypoints: wall time per call and the host-side profile."""
import cProfile, importlib, os, pstats, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "cs566-project-lightglue_b200"
W = importlib.import_module(PKG + ".weights"); M = importlib.import_module(PKG + ".matcher")
conf = {"filter_threshold": 0.1, "depth_confidence": 0.95, "width_confidence": 0.95}
sd = W.make_state_dict(conf, seed=0, adaptive_fixture=True)
model = M.LightGlue(conf); model.load_state_dict(sd); model = model.cuda()
plain = M.LightGlue({"filter_threshold": 0.1}); plain.load_state_dict(sd); plain = plain.cuda()
data = W.make_inputs(1, 2048, 2048, seed=1, correlated=0.7)
data = {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in data.items()}


def wall(m, n=30):
    for _ in range(5):
        m(data)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n):
        m(data)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


print(f"adaptive  : {wall(model):.3f} ms per pair, exit layer {model(data)['exit_layer']}")
print(f"full depth: {wall(plain):.3f} ms per pair")
pr = cProfile.Profile(); pr.enable()
for _ in range(20):
    model(data)
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
