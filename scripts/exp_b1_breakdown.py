"""One pair per call at 2048 keypoints: where the 1.5 ms go -- pos-enc / load, the layer stack (one CUDA graph), the
assignment stage + outputs -- as GPU time (events) and as wall time of each part run alone in a loop."""
import importlib, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "cs566-project-lightglue_b200"
W = importlib.import_module(PKG + ".weights"); M = importlib.import_module(PKG + ".matcher"); C = importlib.import_module(PKG + "._cabi")
conf = {"filter_threshold": 0.1}
sd = W.make_state_dict(conf, seed=0)
model = M.LightGlue(conf); model.load_state_dict(sd); model = model.cuda()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
data = W.make_inputs(1, N, N, seed=1, correlated=0.7)
data = {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in data.items()}
for _ in range(5):
    model(data)
eng = model._engine(1, N, N, torch.device("cuda"), cached=True)
pe = model.posenc
wc = pe.condition_modulation.weight.reshape(-1)
size = data["view0"]["image_size"].float().contiguous()
k0, k1 = data["keypoints0"].contiguous(), data["keypoints1"].contiguous()
d0, d1 = data["descriptors0"].contiguous(), data["descriptors1"].contiguous()


def t(fn, reps=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    host = (time.perf_counter() - w0) / reps * 1e3  # time to ENQUEUE (the loop does not wait for the GPU)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, host


def pre():
    cs0 = C.posenc(k0, size, pe.Wr.weight, wc, pe.condition_modulation.bias, float(N))
    cs1 = C.posenc(k1, size, pe.Wr.weight, wc, pe.condition_modulation.bias, float(N))
    eng.load(d0, d1, cs0, cs1)


def layers():  # (reloaded every time: the stack applied again and again to its own output diverges)
    pre()
    eng.run_layers(9)


def post():
    r0, r1 = eng.rows()
    model._assign(8, eng, r0, r1, None, None, th=0.1)


with C.pinned_stream():
    for name, fn in (("pos-enc + load", pre), ("pos-enc + load + 9 layers (one graph replay)", layers), ("rows + assignment + filter", post)):
        g, h = t(fn)
        print(f"{name:46s} gpu-paced {g:.3f} ms per call   host enqueue {h:.3f} ms per call")
g, h = t(lambda: model(data))
print(f"{'whole forward':46s} gpu-paced {g:.3f} ms per call   host enqueue {h:.3f} ms per call")
