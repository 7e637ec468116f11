"""GPU end-to-end parity: the drop-in matcher against the golden reference outputs and the oracle."""
import numpy as np
import pytest
import torch

from conftest import load_golden, pkg
from oracle import lightglue_oracle as O

pytestmark = pytest.mark.gpu
W = pkg("weights")

BF16_TOL = 2e-2  # north_star: bf16 path, abs on log_assignment


def _cuda(data):
    return {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in data.items()}


def _model(conf, sd):
    M = pkg("matcher")
    model = M.LightGlue(conf)
    missing, unexpected = model.load_state_dict(sd, strict=True)
    return model.cuda()


def _golden_inputs(meta):
    conf = meta["conf"]
    sd = W.make_state_dict(conf, seed=meta["weight_seed"], adaptive_fixture=meta["adaptive_fixture"])
    data = W.make_inputs(meta["batch"], meta["m"], meta["n"], d_in=conf.get("input_dim", 256),
                         seed=meta["data_seed"], image_size=tuple(meta["image_size"]), correlated=meta["correlated"])
    return conf, sd, data


@pytest.mark.parametrize("name", ["sp_b1_192", "sp_b2_160x224", "disk_b1_130x97", "sp_b1_1024"])
def test_forward_vs_golden(name):
    meta, g = load_golden("fwd_" + name)
    conf, sd, data = _golden_inputs(meta)
    pred = _model(conf, sd)(_cuda(data))
    la = pred["log_assignment"].cpu().numpy()
    assert list(la.shape) == meta["la_shape"]
    if "log_assignment" in g:
        err = np.abs(la - g["log_assignment"]).max()
    else:
        err = np.abs(la[:, ::17, ::19] - g["log_assignment_sample"]).max()
    agree0 = (pred["matches0"].cpu().numpy() == g["matches0"]).mean()
    agree1 = (pred["matches1"].cpu().numpy() == g["matches1"]).mean()
    print(f"{name}: max|dlog_assignment|={err:.3e} match agreement {agree0:.4f}/{agree1:.4f}")
    assert err < BF16_TOL, err
    assert agree0 >= 0.995 and agree1 >= 0.995  # measured 1.000
    for k in ("matches0", "matches1", "prune0", "prune1"):
        assert pred[k].cpu().numpy().dtype == g[k].dtype and pred[k].shape == g[k].shape, k
    assert float(pred["prune0"][0, 0]) == 9.0
    assert pred["ref_descriptors0"].shape == (meta["batch"], 1, meta["m"], 256)


FP32_TOL = 1e-4  # north_star: fp32 path, abs on log_assignment


@pytest.mark.parametrize("name", ["sp_b1_192", "sp_b2_160x224", "disk_b1_130x97", "sp_b1_1024"])
def test_forward_fp32_path_vs_golden(name):
    """b200.precision = fp32 (csrc/lg_fp32.cu + TF32x3 assignment): log_assignment within 1e-4 of the reference's fp32
    run, match indices identical up to near-ties."""
    meta, g = load_golden("fwd_" + name)
    conf, sd, data = _golden_inputs(meta)
    pred = _model({**conf, "b200": {"precision": "fp32"}}, sd)(_cuda(data))
    la = pred["log_assignment"].cpu().numpy()
    assert list(la.shape) == meta["la_shape"]
    if "log_assignment" in g:
        err = np.abs(la - g["log_assignment"]).max()
    else:
        err = np.abs(la[:, ::17, ::19] - g["log_assignment_sample"]).max()
    agree0 = (pred["matches0"].cpu().numpy() == g["matches0"]).mean()
    serr = np.abs(pred["matching_scores0"].cpu().numpy() - g["matching_scores0"]).max()
    print(f"fp32 {name}: max|dlog_assignment|={err:.3e} match agreement {agree0:.4f} max|dscore|={serr:.2e}")
    assert err < FP32_TOL, err
    assert agree0 > 0.999 and serr < 1e-4
    assert pred["ref_descriptors0"].shape == (meta["batch"], 1, meta["m"], 256)


def test_forward_fp32_path_varlen_vs_golden():
    meta, g = load_golden("varlen_disk_b3_256")
    conf = {**meta["conf"], "b200": {"precision": "fp32"}}
    sd = W.make_state_dict(meta["conf"], seed=meta["weight_seed"])
    data = _cuda(W.make_inputs(meta["batch"], meta["npad"], meta["npad"], d_in=128, seed=meta["data_seed"],
                               image_size=tuple(meta["image_size"]), correlated=meta["correlated"]))
    data["num_keypoints0"] = torch.tensor(meta["len0"]).cuda()
    data["num_keypoints1"] = torch.tensor(meta["len1"]).cuda()
    pred = _model(conf, sd)(data)
    la = pred["log_assignment"].cpu().numpy()
    M = N = meta["npad"]
    for p in range(meta["batch"]):
        a, c = meta["len0"][p], meta["len1"][p]
        ref = g[f"log_assignment_{p}"][0]
        assert np.abs(la[p, :a, :c] - ref[:a, :c]).max() < FP32_TOL
        assert np.abs(la[p, :a, N] - ref[:a, c]).max() < FP32_TOL
        assert np.abs(la[p, M, :c] - ref[a, :c]).max() < FP32_TOL
        assert (pred["matches0"][p, :a].cpu().numpy() == g[f"matches0_{p}"][0]).mean() > 0.995


def test_forward_varlen_vs_golden():
    meta, g = load_golden("varlen_disk_b3_256")
    conf = meta["conf"]
    sd = W.make_state_dict(conf, seed=meta["weight_seed"])
    data = W.make_inputs(meta["batch"], meta["npad"], meta["npad"], d_in=128, seed=meta["data_seed"],
                         image_size=tuple(meta["image_size"]), correlated=meta["correlated"])
    data = _cuda(data)
    data["num_keypoints0"] = torch.tensor(meta["len0"]).cuda()
    data["num_keypoints1"] = torch.tensor(meta["len1"]).cuda()
    pred = _model(conf, sd)(data)
    la = pred["log_assignment"].cpu().numpy()
    M = N = meta["npad"]
    for p in range(meta["batch"]):
        a, c = meta["len0"][p], meta["len1"][p]
        ref = g[f"log_assignment_{p}"][0]
        assert np.abs(la[p, :a, :c] - ref[:a, :c]).max() < BF16_TOL
        assert np.abs(la[p, :a, N] - ref[:a, c]).max() < BF16_TOL
        assert np.abs(la[p, M, :c] - ref[a, :c]).max() < BF16_TOL
        m0 = pred["matches0"][p, :a].cpu().numpy()
        assert (m0 == g[f"matches0_{p}"][0]).mean() >= 0.995  # measured 1.000
        assert (pred["matches0"][p, a:] == -1).all()


def test_forward_2048_vs_oracle():
    """BASELINE config-2 shape (one pair of it): 2048 keypoints, SuperPoint-shaped."""
    conf = {"filter_threshold": 0.1}
    sd = W.make_state_dict(conf, seed=0)
    data = W.make_inputs(1, 2048, 2048, seed=1, correlated=0.7)
    ref = O.forward(sd, conf, data)
    pred = _model(conf, sd)(_cuda(data))
    err = (pred["log_assignment"].cpu() - ref["log_assignment"]).abs().max().item()
    agree = (pred["matches0"].cpu() == ref["matches0"]).float().mean().item()
    print(f"2048: max|dlog_assignment|={err:.3e} match agreement {agree:.4f}")
    assert err < BF16_TOL and agree >= 0.995  # measured 1.000


def test_forward_fp32_path_2048_vs_oracle():
    """fp32 path at the BASELINE keypoint count (one pair of 2048 x 2048): 1e-4 on log_assignment."""
    conf = {"filter_threshold": 0.1}
    sd = W.make_state_dict(conf, seed=0)
    data = W.make_inputs(1, 2048, 2048, seed=1, correlated=0.7)
    ref = O.forward(sd, conf, data)
    pred = _model({**conf, "b200": {"precision": "fp32"}}, sd)(_cuda(data))
    err = (pred["log_assignment"].cpu() - ref["log_assignment"]).abs().max().item()
    agree = (pred["matches0"].cpu() == ref["matches0"]).float().mean().item()
    print(f"fp32 2048: max|dlog_assignment|={err:.3e} match agreement {agree:.4f}")
    assert err < FP32_TOL and agree > 0.999


@pytest.mark.parametrize("precision,tol", [("bf16", BF16_TOL), ("fp32", 1e-4)])
def test_forward_add_scale_ori_vs_oracle(precision, tol):
    """add_scale_ori = True (lightglue.py:458-476): 4-d positional input [x, y, scale, ori], scales given as (B,N),
    orientations as (B,N,1) -- both layouts the reference accepts."""
    conf = {"filter_threshold": 0.1, "add_scale_ori": True}
    sd = W.make_state_dict(conf, seed=0)
    assert sd["posenc.Wr.weight"].shape == (32, 4)
    data = W.make_inputs(2, 230, 180, seed=6, image_size=(640, 480), correlated=0.7)
    g = torch.Generator().manual_seed(7)
    data["scales0"], data["scales1"] = torch.rand(2, 230, generator=g) * 3 + 0.5, torch.rand(2, 180, generator=g) * 3 + 0.5
    data["oris0"] = (torch.rand(2, 230, 1, generator=g) * 2 - 1) * 3.14159
    data["oris1"] = (torch.rand(2, 180, 1, generator=g) * 2 - 1) * 3.14159
    ref = O.forward(sd, conf, data)
    pred = _model({**conf, "b200": {"precision": precision}}, sd)(_cuda(data))
    err = (pred["log_assignment"].cpu() - ref["log_assignment"]).abs().max().item()
    print(f"add_scale_ori {precision}: max|dlog_assignment|={err:.3e}")
    assert err < tol
    assert (pred["matches0"].cpu() == ref["matches0"]).float().mean() >= 0.995


def test_forward_edge_shapes_vs_oracle():
    """SURVEY 8(c) edge fixtures: a single keypoint in one image, keypoint counts that are not multiples of the tile."""
    conf = {"filter_threshold": 0.0}
    sd = W.make_state_dict(conf, seed=0)
    model = _model(conf, sd)
    for (m, n) in ((129, 1), (1, 1), (257, 383)):
        data = W.make_inputs(1, m, n, seed=5, image_size=(640, 480), correlated=0.5)
        ref = O.forward(sd, conf, data)
        pred = model(_cuda(data))
        assert pred["log_assignment"].shape == ref["log_assignment"].shape == (1, m + 1, n + 1)
        assert (pred["log_assignment"].cpu() - ref["log_assignment"]).abs().max() < BF16_TOL
        assert (pred["matches0"].cpu() == ref["matches0"]).float().mean() >= 0.995
        assert (pred["matching_scores1"].cpu() - ref["matching_scores1"]).abs().max() < BF16_TOL


def test_forward_fully_padded_image_in_batch():
    """A pair whose image has NO valid keypoints inside a variable-length batch: no matches for it, the other pairs
    are what they are without it (the reference cannot run this case: max() over an empty dimension raises)."""
    conf = {"filter_threshold": 0.1}
    sd = W.make_state_dict(conf, seed=0)
    model = _model(conf, sd)
    data = _cuda(W.make_inputs(3, 200, 160, seed=9, image_size=(640, 480), correlated=0.7))
    data["num_keypoints0"] = torch.tensor([200, 0, 150]).cuda()
    data["num_keypoints1"] = torch.tensor([160, 160, 0]).cuda()
    pred = model(data)
    torch.cuda.synchronize()
    assert (pred["matches0"][1] == -1).all() and (pred["matches1"][1] == -1).all()
    assert (pred["matches0"][2] == -1).all() and (pred["matches1"][2] == -1).all()
    assert (pred["matching_scores0"][1:] == 0).all()
    one = model(_slice_pair(data, 0))
    assert (one["matches0"] == pred["matches0"][0:1]).float().mean() > 0.995
    assert torch.isfinite(pred["log_assignment"][0, :200, :160]).all()


def test_registry_alias_and_errors():
    import importlib

    mod = importlib.import_module("b200glue.matcher")
    assert mod.__main_model__ is pkg("matcher").LightGlue
    model = mod.__main_model__({"filter_threshold": 0.1, "some_unknown_key": 1})
    assert model.conf.filter_threshold == 0.1 and model.conf.n_layers == 9
    with pytest.raises(AssertionError):
        model({"keypoints0": torch.zeros(1, 4, 2)})
    with pytest.raises(RuntimeError):  # CPU tensors: no fallback
        d = W.make_inputs(1, 8, 8)
        model(d)


# The adaptive fixtures (exit layer, prune counters, assignment over the kept points) are checked in
# tests/test_scale_parity_gpu.py against the reference's own per-layer score tensors: a keep decision may differ from
# the reference's only for a token within a stated margin of its threshold, and which branch ran is recorded.


def _slice_pair(data, b):
    return {k: ({kk: vv[b:b + 1] for kk, vv in v.items()} if isinstance(v, dict) else v[b:b + 1]) for k, v in data.items()}


@pytest.mark.parametrize("depth,width", [(0.95, 0.95), (-1, 0.95), (0.95, -1)])
def test_adaptive_batched_vs_single_and_oracle(depth, width):
    """A batch of DIFFERENT pairs (and one with fewer valid keypoints): every pair must get the exit layer, prune
    counters and matches it gets when run alone through the same kernels, and pair 0 those of the oracle."""
    conf = {"filter_threshold": 0.1, "depth_confidence": depth, "width_confidence": width}
    sd = W.make_state_dict(conf, seed=0, adaptive_fixture=True)
    data = W.make_inputs(4, 384, 300, seed=11, image_size=(640, 480), correlated=0.7)
    model = _model(conf, sd)
    pred = model(_cuda(data))
    exits = pred["exit_layer"].tolist()
    kept = [tuple(pred["log_assignment"][b].shape[1:]) for b in range(4)]
    print(f"depth={depth} width={width}: exit layers {exits}, kept {kept}")
    for b in range(4):
        one = model(_cuda(_slice_pair(data, b)))
        assert int(one["exit_layer"]) == exits[b]
        assert torch.equal(one["prune0"], pred["prune0"][b:b + 1]) and torch.equal(one["prune1"], pred["prune1"][b:b + 1])
        assert one["log_assignment"].shape == pred["log_assignment"][b].shape
        assert (one["log_assignment"] - pred["log_assignment"][b]).abs().max() < 1e-3
        assert (one["matches0"] == pred["matches0"][b:b + 1]).float().mean() > 0.995
        assert (one["matching_scores1"] - pred["matching_scores1"][b:b + 1]).abs().max() < 1e-3
    ref = O.forward(sd, conf, _slice_pair(data, 0))
    la = pred["log_assignment"][0].cpu()
    if la.shape == ref["log_assignment"].shape:
        assert (la - ref["log_assignment"]).abs().max() < BF16_TOL
        assert (pred["matches0"][0:1].cpu() == ref["matches0"]).float().mean() >= 0.995
    assert (pred["prune0"][0:1].cpu() == ref["prune0"]).float().mean() > 0.99


def test_adaptive_batched_varlen_self_consistent():
    """Adaptive + variable keypoint counts in one batch: pair b equals the same pair run alone with the same padding."""
    conf = {"filter_threshold": 0.1, "depth_confidence": 0.95, "width_confidence": 0.95}
    sd = W.make_state_dict(conf, seed=0, adaptive_fixture=True)
    data = _cuda(W.make_inputs(3, 320, 260, seed=12, image_size=(640, 480), correlated=0.7))
    lens0, lens1 = [320, 200, 131], [260, 260, 97]
    data["num_keypoints0"], data["num_keypoints1"] = torch.tensor(lens0).cuda(), torch.tensor(lens1).cuda()
    model = _model(conf, sd)
    pred = model(data)
    for b in range(3):
        one_data = _slice_pair({k: v for k, v in data.items()}, b)
        one = model(one_data)
        assert one["exit_layer"].tolist() == pred["exit_layer"][b:b + 1].tolist()
        assert torch.equal(one["prune0"], pred["prune0"][b:b + 1])
        assert one["log_assignment"][0].shape == pred["log_assignment"][b].shape
        assert (one["matches0"] == pred["matches0"][b:b + 1]).float().mean() > 0.995
        assert (pred["matches0"][b, lens0[b]:] == -1).all() and (pred["prune0"][b, lens0[b]:] == 0).all()


def test_cuda_graph_replay_matches_eager():
    """Same shape called repeatedly: eager first call, graph capture on the second, replay afterwards.
    Every call must give bit-identical results to a graph-free model on the same inputs."""
    conf = {"filter_threshold": 0.1}
    sd = W.make_state_dict(conf, seed=0)
    eager = _model({**conf, "b200": {"cuda_graph": False}}, sd)
    graphed = _model(conf, sd)
    for seed in (3, 4, 5, 3):
        data = _cuda(W.make_inputs(2, 300, 200, seed=seed, correlated=0.6))
        a, b = eager(data), graphed(data)
        for k in ("log_assignment", "matches0", "matches1", "matching_scores0", "ref_descriptors1"):
            assert torch.equal(a[k], b[k]), (seed, k)
    # variable-length batches go through their own captured graph with static length buffers
    for seed, lens in ((6, ([300, 150], [200, 31])), (7, ([10, 299], [200, 200])), (8, ([300, 300], [1, 77]))):
        data = _cuda(W.make_inputs(2, 300, 200, seed=seed, correlated=0.6))
        data["num_keypoints0"], data["num_keypoints1"] = torch.tensor(lens[0]).cuda(), torch.tensor(lens[1]).cuda()
        a, b = eager(data), graphed(data)
        assert torch.equal(a["log_assignment"], b["log_assignment"]) and torch.equal(a["matches0"], b["matches0"])


def test_host_pipeline_matches_direct_calls():
    """HostPipeline (copies on side streams, double-buffered slots) returns, batch by batch and in order, exactly
    what model(batch.cuda()) returns -- including when the host buffers are rewritten right after submission."""
    P = pkg("pipeline")
    conf = {"filter_threshold": 0.1}
    model = _model(conf, W.make_state_dict(conf, seed=0))
    batches = [W.make_inputs(2, 260, 190, seed=s, correlated=0.6) for s in (11, 12, 13, 14, 15)]
    want = [{k: model(_cuda(b))[k].cpu() for k in P.RESULT_KEYS} for b in batches]
    pinned = [{k: ({kk: vv.pin_memory() for kk, vv in v.items()} if isinstance(v, dict) else v.pin_memory())
               for k, v in b.items()} for b in batches]
    got = [{k: v.clone() for k, v in pred.items()} for pred in P.HostPipeline(model, depth=2).run(iter(pinned))]
    assert len(got) == len(want)
    for i, (a, b) in enumerate(zip(want, got)):
        for k in P.RESULT_KEYS:
            assert torch.equal(a[k], b[k]), (i, k)


def test_host_pipeline_shape_change_and_result_lifetime():
    """A run whose batches change shape (groups of pack_pairs pad to different lengths): the batch in flight is handed
    out before the pinned slots are re-allocated, nothing is lost or returned uninitialised; and a yielded result
    stays valid, WITHOUT a clone, until `depth` more batches have been yielded."""
    P = pkg("pipeline")
    conf = {"filter_threshold": 0.1}
    model = _model(conf, W.make_state_dict(conf, seed=0))
    shapes = [(260, 190), (260, 190), (400, 300), (400, 300), (400, 300), (130, 260), (260, 190)]
    batches = [W.make_inputs(2, m, n, seed=20 + i, correlated=0.6) for i, (m, n) in enumerate(shapes)]
    want = [{k: model(_cuda(b))[k].cpu() for k in P.RESULT_KEYS} for b in batches]
    pinned = [{k: ({kk: vv.pin_memory() for kk, vv in v.items()} if isinstance(v, dict) else v.pin_memory())
               for k, v in b.items()} for b in batches]
    depth, held, checked = 2, [], 0
    for i, pred in enumerate(P.HostPipeline(model, depth=depth).run(iter(pinned))):
        for k in P.RESULT_KEYS:
            assert torch.equal(want[i][k], pred[k]), (i, k)
        held.append((i, pred))
        for j, old in held[-depth:]:  # the last `depth` results are still intact (same shape group or not)
            for k in P.RESULT_KEYS:
                assert torch.equal(want[j][k], old[k]), ("held", j, k)
                checked += 1
    assert len(held) == len(batches) and checked > 0


@pytest.mark.parametrize("precision,la_tol", [("bf16", 1e-4), ("fp32", 1e-5)])
def test_packed_pairs_reproduce_one_pair_per_call(precision, la_tol):
    """pipeline.pack_pairs + conf b200.condition = "valid": pairs with unequal keypoint counts in ONE padded batch give,
    pair by pair, what one call per pair gives (the only mode of the reference's exporter,
    utils/export_predictions.py:32-79; the reference conditions the positional encoding on the count,
    lightglue.py:490-491: a common phase of q and k, which cancels in exact arithmetic but moves every sin / cos by
    rounding, so a plainly padded batch agrees only to the path's noise level) -- and with it the oracle's one-pair forward."""
    P = pkg("pipeline")
    conf = {"filter_threshold": 0.1, "b200": {"condition": "valid", "precision": precision}}
    sd = W.make_state_dict(conf, seed=0)
    model = _model(conf, sd)
    shapes = [(300, 222), (517, 640), (700, 511), (129, 128)]
    pairs = [W.make_inputs(1, m, n, seed=40 + i, correlated=0.7) for i, (m, n) in enumerate(shapes)]
    pairs[1].pop("view0"), pairs[1].pop("view1")  # the size=None branch must not see the padding either
    batch = P.pack_pairs([_cuda(p) for p in pairs])
    assert batch["keypoints0"].shape == (4, 768, 2) and batch["keypoints1"].shape == (4, 640, 2)
    pred = model(batch)
    per_pair = P.unpack_pred(pred, batch)
    for i, (p, got) in enumerate(zip(pairs, per_pair)):
        single = model(_cuda(p))
        for k in ("matches0", "matches1"):
            assert torch.equal(single[k], got[k]), (i, k)
        for k in ("matching_scores0", "matching_scores1"):
            assert (single[k] - got[k]).abs().max() < 1e-5, (i, k)
        la_s, la_b = single["log_assignment"], got["log_assignment"]
        assert la_s.shape == la_b.shape
        assert (la_s - la_b).abs().max() < la_tol, i
        want = O.forward(sd, conf, p)  # the reference's own one-pair forward (oracle pinned to it by the goldens)
        err = (want["log_assignment"] - la_b.cpu()).abs().max().item()
        assert err < (BF16_TOL if precision == "bf16" else 1e-4), (i, err)
        assert (want["matches0"] == got["matches0"].cpu()).float().mean() >= 0.995


def test_packed_pairs_at_export_scale():
    """16 pairs of 1024..2048 keypoints packed to 2048 (the exporter workload): many work items per persistent CTA with
    skipped blocks in between -- the shape that dead-locked the attention kernel before the P-ready barriers alternated."""
    P = pkg("pipeline")
    conf = {"filter_threshold": 0.1, "b200": {"condition": "valid"}}
    model = _model(conf, W.make_state_dict(conf, seed=0))
    rng = np.random.default_rng(7)
    counts = rng.integers(1024, 2049, (16, 2))
    pairs = [_cuda(W.make_inputs(1, int(m), int(n), seed=100 + i, correlated=0.7)) for i, (m, n) in enumerate(counts)]
    batch = P.pack_pairs(pairs, length0=2048, length1=2048)
    for _ in range(2):  # eager call, then the captured graph
        per_pair = P.unpack_pred(model(batch), batch)
    for i in (0, 9, 15):
        single = model(pairs[i])
        assert torch.equal(single["matches0"], per_pair[i]["matches0"])
        assert (single["log_assignment"] - per_pair[i]["log_assignment"]).abs().max() < 1e-4


def test_forward_large_batch_equals_one_pair_calls():
    """16 x 2048 x 2048 (1024 attention work items on 148 persistent CTAs, 7 per CTA) against the same pairs one per
    call (64 items: one per CTA).  Caught a premature read of the attention accumulator at item boundaries (parity wait
    on a barrier two completions behind) that only showed with several items per CTA."""
    conf = {"filter_threshold": 0.1}
    model = _model(conf, W.make_state_dict(conf, seed=0))
    data = _cuda(W.make_inputs(16, 2048, 2048, seed=3, correlated=0.7))
    for _ in range(3):  # eager, capture, replay
        pred = model(data)
        la = pred["log_assignment"].clone()
        for b in (0, 5, 10, 15):
            single = model(_slice_pair(data, b))
            assert (single["log_assignment"][0] - la[b]).abs().max() < 1e-4, b
            assert torch.equal(single["matches0"][0], pred["matches0"][b])
    # and against the oracle, not only against itself: two pairs of the batch through the CPU restatement
    sd = W.make_state_dict(conf, seed=0)
    host = W.make_inputs(16, 2048, 2048, seed=3, correlated=0.7)
    for b in (0, 15):
        ref = O.forward(sd, conf, _slice_pair(host, b))
        err = (la[b].cpu() - ref["log_assignment"][0]).abs().max().item()
        agree = (pred["matches0"][b].cpu() == ref["matches0"][0]).float().mean().item()
        print(f"16 x 2048, pair {b}: max|dlog_assignment|={err:.3e} match agreement {agree:.4f}")
        assert err < BF16_TOL and agree >= 0.995


def test_adaptive_batched_at_scale_vs_single():
    """Adaptive depth + width on 16 pairs x 2048 keypoints (several work items per persistent CTA in every kernel, pruned
    lengths differ per image): pairs 0, 7, 15 must match the same pair run alone through the B = 1 adaptive path."""
    conf = {"filter_threshold": 0.1, "depth_confidence": 0.95, "width_confidence": 0.95}
    sd = W.make_state_dict(conf, seed=0, adaptive_fixture=True)
    data = _cuda(W.make_inputs(16, 2048, 2048, seed=5, correlated=0.7))
    model = _model(conf, sd)
    pred = model(data)
    for b in (0, 7, 15):
        one = model(_slice_pair(data, b))
        assert int(one["exit_layer"]) == int(pred["exit_layer"][b])
        assert torch.equal(one["prune0"], pred["prune0"][b:b + 1]) and torch.equal(one["prune1"], pred["prune1"][b:b + 1])
        assert one["log_assignment"].shape == pred["log_assignment"][b].shape
        assert (one["log_assignment"] - pred["log_assignment"][b]).abs().max() < 1e-3
        assert (one["matches0"] == pred["matches0"][b:b + 1]).float().mean() > 0.995


def test_forward_soak_random_lengths():
    """Timing-dependent faults need more than one shape: ten padded batches with random sizes and valid counts (some
    images with very few or no points); two pairs of each batch against the same pair run alone with the same padding
    and counts (same condition, same arithmetic, one work item per CTA)."""
    conf = {"filter_threshold": 0.1}
    model = _model(conf, W.make_state_dict(conf, seed=0))
    rng = np.random.default_rng(123)
    for it in range(10):
        B = int(rng.choice([6, 12, 20]))
        m, n = int(rng.choice([640, 1024, 1536, 2048])), int(rng.choice([640, 1024, 1536, 2048]))
        data = _cuda(W.make_inputs(B, m, n, seed=200 + it, correlated=0.7))
        l0, l1 = rng.integers(m // 3, m + 1, B), rng.integers(n // 3, n + 1, B)
        l0[rng.integers(B)] = int(rng.choice([0, 1, 5]))
        l1[rng.integers(B)] = int(rng.choice([0, 1, 127]))
        data["num_keypoints0"], data["num_keypoints1"] = torch.tensor(l0).cuda(), torch.tensor(l1).cuda()
        pred = model(data)
        la = pred["log_assignment"]
        for b in (int(rng.integers(B)), B - 1):
            one = model(_slice_pair(data, b))
            fin = torch.isfinite(one["log_assignment"][0])
            assert torch.equal(fin, torch.isfinite(la[b])), (it, b)
            assert (one["log_assignment"][0][fin] - la[b][fin]).abs().max() < 1e-4, (it, b, B, m, n)
            assert torch.equal(one["matches0"][0], pred["matches0"][b]), (it, b)


def test_adaptive_batched_varlen_at_scale():
    """Adaptive depth + width AND per-image valid counts on 12 pairs padded to 2048: pairs 0, 5, 11 against the same
    pair run alone with the same padding and counts."""
    conf = {"filter_threshold": 0.1, "depth_confidence": 0.95, "width_confidence": 0.95}
    sd = W.make_state_dict(conf, seed=0, adaptive_fixture=True)
    data = _cuda(W.make_inputs(12, 2048, 2048, seed=6, correlated=0.7))
    rng = np.random.default_rng(5)
    l0, l1 = rng.integers(700, 2049, 12), rng.integers(700, 2049, 12)
    data["num_keypoints0"], data["num_keypoints1"] = torch.tensor(l0).cuda(), torch.tensor(l1).cuda()
    model = _model(conf, sd)
    pred = model(data)
    for b in (0, 5, 11):
        one = model(_slice_pair(data, b))
        assert one["exit_layer"].tolist() == pred["exit_layer"][b:b + 1].tolist()
        assert torch.equal(one["prune0"], pred["prune0"][b:b + 1]) and torch.equal(one["prune1"], pred["prune1"][b:b + 1])
        assert one["log_assignment"][0].shape == pred["log_assignment"][b].shape
        assert (one["matches0"] == pred["matches0"][b:b + 1]).float().mean() > 0.995
        assert (pred["matches0"][b, int(l0[b]):] == -1).all()
