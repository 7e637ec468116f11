"""BASELINE configs 3 and 4 on one GPU; bench.py carries config 2.
  config 3: adaptive LightGlue, depth_confidence = width_confidence = 0.95, 32 pairs x 2048 kpts, adaptive fixture
            weights (random init never exits): pairs/s, exit-layer histogram, kept-token counts.
  config 4: DISK-shaped 128-d descriptors (input_proj), padded to 4096 kpts, per-image valid counts U{2048..4096}.
Prints one JSON line per config (device-resident inputs, CUDA events, 3 warm-ups)."""
import collections, importlib, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "cs566-project-lightglue_b200"
W = importlib.import_module(PKG + ".weights")
M = importlib.import_module(PKG + ".matcher")


def cuda(data):
    return {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in data.items()}


def timed(model, data, steps, warmup=10):  # a fresh box needs ~10 forwards to ramp its clocks / page in
    for _ in range(warmup):
        pred = model(data)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        pred = model(data)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, pred


def config3(B=32, N=2048, steps=5):
    conf = {"filter_threshold": 0.1, "depth_confidence": 0.95, "width_confidence": 0.95}
    sd = W.make_state_dict(conf, seed=0, adaptive_fixture=True)
    model = M.LightGlue(conf)
    model.load_state_dict(sd)
    model = model.cuda()
    data = cuda(W.make_inputs(B, N, N, seed=1, correlated=0.7))
    ms, pred = timed(model, data, steps)
    exits = collections.Counter(pred["exit_layer"].tolist())
    kept = [tuple(int(x) - 1 for x in pred["log_assignment"][b].shape[1:]) for b in range(B)]
    one = cuda(W.make_inputs(1, N, N, seed=1, correlated=0.7))
    ms1, _ = timed(model, one, steps)
    full = M.LightGlue({"filter_threshold": 0.1})
    full.load_state_dict(sd)
    ms_full, _ = timed(full.cuda(), data, steps)
    print(json.dumps({"config": "3: adaptive depth=width=0.95, %d pairs x %d kpts, adaptive-fixture weights" % (B, N),
                      "pairs_per_s": B / ms * 1e3, "ms_per_batch": ms, "exit_layer_histogram": dict(sorted(exits.items())),
                      "kept0_min_mean_max": [min(k[0] for k in kept), float(np.mean([k[0] for k in kept])), max(k[0] for k in kept)],
                      "kept1_min_mean_max": [min(k[1] for k in kept), float(np.mean([k[1] for k in kept])), max(k[1] for k in kept)],
                      "one_pair_per_call_pairs_per_s": 1e3 / ms1,
                      "non_adaptive_same_weights_pairs_per_s": B / ms_full * 1e3}))


def config4(B=8, NPAD=4096, steps=5):
    # filter_threshold 0: a random-init network never passes 0.1 at 4096 points, and a configuration without matches
    # says nothing about the indices (parity at this size: tests/test_scale_parity_gpu.py::test_forward_varlen_4096_vs_golden)
    conf = {"filter_threshold": 0.0, "input_dim": 128}
    sd = W.make_state_dict(conf, seed=0)
    model = M.LightGlue(conf)
    model.load_state_dict(sd)
    model = model.cuda()
    data = cuda(W.make_inputs(B, NPAD, NPAD, d_in=128, seed=2, correlated=0.7))
    rng = np.random.default_rng(4)
    l0, l1 = rng.integers(2048, NPAD + 1, B), rng.integers(2048, NPAD + 1, B)
    data["num_keypoints0"], data["num_keypoints1"] = torch.tensor(l0).cuda(), torch.tensor(l1).cuda()
    ms, pred = timed(model, data, steps)
    print(json.dumps({"config": "4: DISK-shaped 128-d descriptors, %d pairs padded to %d kpts, valid counts U{2048..%d}" % (B, NPAD, NPAD),
                      "pairs_per_s": B / ms * 1e3, "ms_per_batch": ms, "valid0": l0.tolist(), "valid1": l1.tolist(),
                      "matches_per_pair_mean": float((pred["matches0"] >= 0).sum(1).float().mean())}))


def fp32_path(N=2048, steps=3):
    conf = {"filter_threshold": 0.1, "b200": {"precision": "fp32"}}
    sd = W.make_state_dict(conf, seed=0)
    model = M.LightGlue(conf)
    model.load_state_dict(sd)
    model = model.cuda()
    out = {"config": "fp32 path (b200.precision=fp32), %d kpts" % N}
    for B in (1, 4, 32):
        ms, _ = timed(model, cuda(W.make_inputs(B, N, N, seed=1, correlated=0.7)), steps, warmup=1)
        out["pairs_per_s_B%d" % B] = B / ms * 1e3
    print(json.dumps(out))


def exporter_loop(pairs_total=64, group=16, lo=1024, hi=2048):
    """The exporter's workload (utils/export_predictions.py:32-79): single pairs with unequal keypoint counts.
    (a) one call per pair, as the reference drives the matcher; (b) pipeline.pack_pairs in groups + b200.condition=valid
    (same results pair by pair, tests/test_forward_gpu.py::test_packed_pairs_reproduce_one_pair_per_call)."""
    P = importlib.import_module(PKG + ".pipeline")
    conf = {"filter_threshold": 0.1, "b200": {"condition": "valid"}}
    model = M.LightGlue(conf)
    model.load_state_dict(W.make_state_dict(conf, seed=0))
    model = model.cuda()
    rng = np.random.default_rng(7)
    counts = rng.integers(lo, hi + 1, (pairs_total, 2))
    pairs = [cuda(W.make_inputs(1, int(m), int(n), seed=100 + i, correlated=0.7)) for i, (m, n) in enumerate(counts)]

    def one_by_one():
        return [model(p)["matches0"] for p in pairs]

    def packed():
        out = []
        for g in range(0, pairs_total, group):
            batch = P.pack_pairs(pairs[g:g + group], length0=hi, length1=hi)
            out += [d["matches0"] for d in P.unpack_pred(model(batch), batch)]
        return out

    sorted_model = M.LightGlue({"filter_threshold": 0.1, "b200": {"condition": "valid", "engine_cache": 8}})
    sorted_model.load_state_dict(W.make_state_dict(conf, seed=0))
    sorted_model = sorted_model.cuda()

    def packed_sorted():  # similar sizes together, each group padded to its own longest pair (multiple of 256)
        out = [None] * pairs_total
        for idx in P.group_by_length(pairs, group):
            batch = P.pack_pairs([pairs[i] for i in idx], multiple=256)
            for i, d in zip(idx, P.unpack_pred(sorted_model(batch), batch)):
                out[i] = d["matches0"]
        return out

    res = {}
    for name, fn in (("one_pair_per_call", one_by_one), ("packed_groups_of_%d" % group, packed),
                     ("packed_sorted_groups_of_%d" % group, packed_sorted)):
        for _ in range(2):
            got = fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        got = fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = pairs_total / e0.elapsed_time(e1) * 1e3
        res.setdefault("_m", []).append(got)
    a, b, c = res.pop("_m")
    same = sum(int(torch.equal(x, y) and torch.equal(x, z)) for x, y, z in zip(a, b, c))
    print(json.dumps({"config": "exporter loop: %d single pairs, counts U{%d..%d} per image" % (pairs_total, lo, hi),
                      "pairs_per_s": res, "pairs_with_identical_matches0": "%d/%d" % (same, pairs_total)}))


if __name__ == "__main__":
    import sys
    if "fp32" in sys.argv[1:]:
        fp32_path()
    elif "exporter" in sys.argv[1:]:
        exporter_loop()
    else:
        config3()
        config4()
        fp32_path()
