"""Where the batched adaptive forward spends its time (32 pairs x 2048): layers eager vs graph, with / without lengths,
and the per-layer scoring / compaction extras."""
import importlib, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "cs566-project-lightglue_b200"
W = importlib.import_module(PKG + ".weights"); M = importlib.import_module(PKG + ".matcher"); C = importlib.import_module(PKG + "._cabi")
conf = {"filter_threshold": 0.1}
sd = W.make_state_dict(conf, seed=0, adaptive_fixture=True)
model = M.LightGlue(conf); model.load_state_dict(sd); model = model.cuda()
B, N = 32, 2048
data = W.make_inputs(B, N, N, seed=1, correlated=0.7)
data = {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in data.items()}
model(data); model(data); model(data)
eng = model._engine(B, N, N, torch.device("cuda"), cached=True)
lens = torch.full((B,), N, dtype=torch.int32, device="cuda")
l18 = torch.full((B,), 1830, dtype=torch.int32, device="cuda")


def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    w0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, (time.perf_counter() - w0) * 1e3 / reps


def layers(n, l0=None, l1=None):
    for i in range(n):
        eng.layer(i, l0, l1)


pe = model.posenc
wc = pe.condition_modulation.weight.reshape(-1)
size = data["view0"]["image_size"].float().contiguous()
cs0 = C.posenc(data["keypoints0"].contiguous(), size, pe.Wr.weight, wc, pe.condition_modulation.bias, float(N))
cs1 = C.posenc(data["keypoints1"].contiguous(), size, pe.Wr.weight, wc, pe.condition_modulation.bias, float(N))
desc0, desc1 = data["descriptors0"].contiguous(), data["descriptors1"].contiguous()


def fresh(n, l0=None, l1=None, graph=False):
    eng.load(desc0, desc1, cs0, cs1)
    if graph:
        eng.run_layers(n)
    else:
        layers(n, l0, l1)


print("load + 6 layers eager, no lens      gpu %.2f ms  wall %.2f ms" % t(lambda: fresh(6)))
print("load + 6 layers eager, lens = 2048  gpu %.2f ms  wall %.2f ms" % t(lambda: fresh(6, lens, lens)))
print("load + 6 layers eager, lens = 1830  gpu %.2f ms  wall %.2f ms" % t(lambda: fresh(6, l18, l18)))
print("load + 9 layers graph               gpu %.2f ms  wall %.2f ms" % t(lambda: fresh(9, graph=True)))
print("load + 9 layers eager               gpu %.2f ms  wall %.2f ms" % t(lambda: fresh(9)))
amodel = M.LightGlue({"filter_threshold": 0.1, "depth_confidence": 0.95, "width_confidence": 0.95})
amodel.load_state_dict(sd); amodel = amodel.cuda()
print("adaptive forward (exit at 5)        gpu %.2f ms  wall %.2f ms" % t(lambda: amodel(data)))
dmodel = M.LightGlue({"filter_threshold": 0.1, "depth_confidence": 0.95})
dmodel.load_state_dict(sd); dmodel = dmodel.cuda()
print("depth-only forward                  gpu %.2f ms  wall %.2f ms" % t(lambda: dmodel(data)))
head = model.log_assignment[0].matchability
tok = model.token_confidence[0].token[0]
ind0 = torch.arange(N, device="cuda")[None].repeat(B, 1)
ind1 = ind0.clone()
pr0 = torch.ones(B, N, dtype=torch.int64, device="cuda")
pr1 = pr0.clone()
counts = torch.empty(B, 4, dtype=torch.int32, device="cuda")
i0n, i1n, p0n, p1n = torch.empty_like(ind0), torch.empty_like(ind1), torch.empty_like(pr0), torch.empty_like(pr1)
l0n, l1n = torch.empty_like(lens), torch.empty_like(lens)


def fused():  # the adaptive step of one layer: lg_score_prune_tiles into the engine's second buffers
    C.score_prune_tiles(eng.x, eng.rot, eng.x_alt, eng.rot_alt(), (ind0, ind1), (i0n, i1n), (pr0, pr1), (p0n, p1n),
                        (lens, lens), (l0n, l1n), counts, tok.weight.reshape(-1), tok.bias, head.weight.reshape(-1),
                        head.bias, 0.85, 0.05, True, B, N, N, eng.tp0, eng.tp1, N, N)


def exit_assign():  # what a layer with exits adds: rows for the logits, final_proj, assignment, filter
    d0, d1 = eng.rows()
    model._assign(5, eng, d0, d1, l18, l18, th=0.1)


print("adaptive step (fused kernel)        gpu %.3f ms  wall %.3f ms" % t(fused))
print("exit assignment (lens = 1830)       gpu %.3f ms  wall %.3f ms" % t(exit_assign))
