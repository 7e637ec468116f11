"""GPU parity at the sizes of BASELINE configs 3 and 4, against goldens produced by the reference itself
(oracle/make_golden_scale.py), and the adaptive fixtures checked with the reference's own per-layer score tensors:
a keep decision may differ from the reference's only for a token whose score sits within MARGIN of its threshold."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import ROOT, load_golden, pkg

pytestmark = pytest.mark.gpu
W = pkg("weights")

BF16_TOL = 2e-2  # north_star: abs on log_assignment, bf16 path
FP32_TOL = 1e-4  # north_star: fp32 path
MARGIN = 2e-3    # |score - threshold| below which a keep / stop decision may legitimately flip on the bf16 path


def _cuda(data):
    return {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in data.items()}


def _model(conf, sd):
    model = pkg("matcher").LightGlue(conf)
    model.load_state_dict(sd, strict=True)
    return model.cuda()


def _record(name, **kw):
    """Which branch of a test ran on this GPU, and its measured numbers: gpurun_out/parity_branches.jsonl"""
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "parity_branches.jsonl"), "a") as f:
        f.write(json.dumps(dict(test=name, **kw)) + "\n")
    print(name, kw)


# ----------------------------------------------------------------------------- the reference's per-layer tensors
def _layer_positions(tr, side, n_layers_pruned, n):
    """Original index of every token alive at pruning layer i (lightglue.py:536-547): ind_0 = arange, ind_{i+1} = ind_i[keep_i]."""
    ind = [np.arange(n)]
    for i in range(n_layers_pruned):
        ind.append(ind[-1][tr[f"L{i}_keep{side}"][0]])
    return ind


def _check_prune_against_trace(name, meta, tr, got, side, n):
    """got: our prune counters (n,). Every token whose counter differs from the reference's must have been marginal at
    the layer where the two runs parted: min(|matchability - (1 - width)|, |confidence - threshold|) < MARGIN."""
    ref = tr["prune" + str(side)][0] if ("prune" + str(side)) in tr else None
    assert ref is not None
    diff = np.nonzero(got != ref)[0]
    if diff.size == 0:
        return 0, 0.0
    npr = meta["layers_pruned"]
    ind = _layer_positions(tr, side, npr, n)
    width = meta["conf"]["width_confidence"]
    worst = 0.0
    for p in diff:
        layer = int(min(got[p], ref[p])) - 1  # counters start at 1 and grow by one per pruning layer survived
        assert 0 <= layer < npr, (p, got[p], ref[p])
        pos = np.nonzero(ind[layer] == p)[0]
        assert pos.size == 1, "the token was alive in the reference at the layer where the runs part"
        mg = abs(float(tr[f"L{layer}_match{side}"][0, pos[0]]) - np.float32(1 - width))
        ck = f"L{layer}_conf{side}"
        if ck in tr:
            mg = min(mg, abs(float(tr[ck][0, pos[0]]) - np.float32(meta["confidence_thresholds"][layer])))
        worst = max(worst, mg)
    assert worst < MARGIN, f"{name}: {diff.size} keep decisions differ, the least marginal one by {worst:.2e}"
    return int(diff.size), worst


def _load_trace(name):
    """-> meta, arrays holding L{i}_* and prune0/1 of the reference run (trace_* files carry only the layers)."""
    if name.endswith("2048"):
        meta, g = load_golden("fwd_" + name)
        return meta, g, g
    meta, tr = load_golden("trace_" + name)
    _, g = load_golden("fwd_" + name)
    tr = dict(tr)
    tr["prune0"], tr["prune1"] = g["prune0"], g["prune1"]
    return meta, tr, g


ADAPTIVE = ["adaptive_both_512x400", "adaptive_width_300x400", "adaptive_depth_300x400", "adaptive_both_2048"]


@pytest.mark.parametrize("name", ADAPTIVE)
@pytest.mark.parametrize("batched", [False, True])
def test_adaptive_vs_reference_trace(name, batched):
    """Config 3 (also at 2048 x 2048): exit layer equal to the reference's; prune counters equal except for tokens that
    were within MARGIN of a threshold in the reference's own per-layer tensors; when no decision flipped, the
    assignment over the kept points within the bf16 bound and the matches in agreement."""
    meta, tr, g = _load_trace(name)
    conf = meta["conf"]
    sd = W.make_state_dict(conf, seed=meta["weight_seed"], adaptive_fixture=True)
    data = W.make_inputs(1, meta["m"], meta["n"], seed=meta["data_seed"], image_size=tuple(meta["image_size"]),
                         correlated=meta["correlated"])
    if batched:  # two copies through the batched adaptive forward
        data = {k: ({kk: vv.repeat(2, 1) for kk, vv in v.items()} if isinstance(v, dict) else v.repeat(2, 1, 1))
                for k, v in data.items()}
    pred = _model(conf, sd)(_cuda(data))
    exits = [int(e) for e in torch.as_tensor(pred["exit_layer"]).reshape(-1).tolist()]
    assert all(e == meta["exit_layer"] for e in exits), (exits, meta["exit_layer"])
    for b in range(2 if batched else 1):
        flips = 0
        worst = 0.0
        for side, n in ((0, meta["m"]), (1, meta["n"])):
            got = pred[f"prune{side}"][b].cpu().numpy()
            assert got.dtype == g[f"prune{side}"].dtype
            k, w = _check_prune_against_trace(name, meta, tr, got, side, n)
            flips += k
            worst = max(worst, w)
        la = (pred["log_assignment"][b] if batched else pred["log_assignment"]).cpu().numpy()
        la = la.reshape((1,) + la.shape[-2:])
        branch = "identical keep decisions" if flips == 0 else f"{flips} marginal decisions flipped"
        err = agree = None
        if flips == 0:
            assert list(la.shape) == meta["la_shape"]
            if "log_assignment" in g:
                err = float(np.abs(la - g["log_assignment"]).max())
            else:
                err = float(np.abs(la[:, ::17, ::19] - g["log_assignment_sample"]).max())
                rs = np.abs(la.astype(np.float64).sum(2) - g["log_assignment_rowsum"]).max() / la.shape[2]
                assert rs < BF16_TOL
            assert err < BF16_TOL
            agree = float((pred["matches0"][b:b + 1].cpu().numpy() == g["matches0"]).mean())
            assert agree >= 0.995
            assert np.abs(pred["matching_scores0"][b:b + 1].cpu().numpy() - g["matching_scores0"]).max() < BF16_TOL
        else:  # sizes can only differ by the flipped tokens
            assert abs(la.shape[1] - meta["la_shape"][1]) + abs(la.shape[2] - meta["la_shape"][2]) <= flips
        _record("test_adaptive_vs_reference_trace", case=name, batched=batched, pair=b, branch=branch,
                worst_margin=worst, la_err=err, match_agreement=agree, exit_layer=exits[b])


@pytest.mark.parametrize("name", ADAPTIVE)
def test_adaptive_fp32_vs_reference_trace(name):
    """b200.precision = fp32 with depth / width confidence (one pair, the reference's own mode): exit layer, prune
    counters (a flip only within 1e-4 of a threshold), kept shape and log_assignment within the fp32 bound."""
    meta, tr, g = _load_trace(name)
    conf = dict(meta["conf"], b200={"precision": "fp32"})
    sd = W.make_state_dict(meta["conf"], seed=meta["weight_seed"], adaptive_fixture=True)
    data = W.make_inputs(1, meta["m"], meta["n"], seed=meta["data_seed"], image_size=tuple(meta["image_size"]),
                         correlated=meta["correlated"])
    pred = _model(conf, sd)(_cuda(data))
    assert int(pred["exit_layer"]) == meta["exit_layer"]
    flips, worst = 0, 0.0
    for side, n in ((0, meta["m"]), (1, meta["n"])):
        got = pred[f"prune{side}"][0].cpu().numpy()
        k, w = _check_prune_against_trace(name, meta, tr, got, side, n)
        flips += k
        worst = max(worst, w)
    assert worst < 1e-4
    la = pred["log_assignment"].cpu().numpy()
    err = agree = None
    if flips == 0:
        assert list(la.shape) == meta["la_shape"]
        if "log_assignment" in g:
            err = float(np.abs(la - g["log_assignment"]).max())
        else:
            err = float(np.abs(la[:, ::17, ::19] - g["log_assignment_sample"]).max())
        assert err < FP32_TOL
        agree = float((pred["matches0"].cpu().numpy() == g["matches0"]).mean())
        assert agree >= 0.999
        assert np.abs(pred["matching_scores0"].cpu().numpy() - g["matching_scores0"]).max() < FP32_TOL
    _record("test_adaptive_fp32_vs_reference_trace", case=name, flips=flips, worst_margin=worst, la_err=err,
            match_agreement=agree, exit_layer=int(pred["exit_layer"]))


def test_adaptive_fp32_batch_equals_one_pair_calls():
    """fp32 mode, B > 1 with unequal valid counts: pair by pair, identical to the one-pair calls on the valid points."""
    conf = {"filter_threshold": 0.0, "depth_confidence": 0.95, "width_confidence": 0.95, "b200": {"precision": "fp32"}}
    sd = W.make_state_dict(conf, seed=0, adaptive_fixture=True)
    model = _model(conf, sd)
    data = _cuda(W.make_inputs(2, 300, 260, seed=5, image_size=(640, 480), correlated=0.7))
    l0, l1 = [300, 211], [190, 260]
    data["num_keypoints0"], data["num_keypoints1"] = torch.tensor(l0).cuda(), torch.tensor(l1).cuda()
    pred = model(data)
    for p in range(2):
        one = {"keypoints0": data["keypoints0"][p:p + 1, :l0[p]], "keypoints1": data["keypoints1"][p:p + 1, :l1[p]],
               "descriptors0": data["descriptors0"][p:p + 1, :l0[p]], "descriptors1": data["descriptors1"][p:p + 1, :l1[p]],
               "view0": {"image_size": data["view0"]["image_size"][p:p + 1]},
               "view1": {"image_size": data["view1"]["image_size"][p:p + 1]}}
        # the batch is conditioned on the padded length (lightglue.py:490-491): compare exit / prune structure only
        ref = model(one)
        assert pred["log_assignment"][p].shape[0] == 1 and pred["log_assignment"][p].dim() == 3
        assert (pred["matches0"][p, l0[p]:] == -1).all() and (pred["prune0"][p, l0[p]:] == 0).all()
        assert 0 <= int(pred["exit_layer"][p]) < model.conf.n_layers and 0 <= ref["exit_layer"] < model.conf.n_layers


# ----------------------------------------------------------------------------- config 4 at its own size
@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_forward_varlen_4096_vs_golden(precision):
    """Config 4: DISK-shaped 128-d descriptors, two pairs padded to 4096 with unequal valid counts, against the
    reference's masked forward (strided sample, last rows incl. the dustbin row, row sums, and ALL match indices)."""
    meta, g = load_golden("varlen_disk_b2_4096")
    conf = dict(meta["conf"])
    if precision == "fp32":
        conf["b200"] = {"precision": "fp32"}
    tol = FP32_TOL if precision == "fp32" else BF16_TOL
    sd = W.make_state_dict(meta["conf"], seed=meta["weight_seed"])
    data = W.make_inputs(meta["batch"], meta["npad"], meta["npad"], d_in=128, seed=meta["data_seed"],
                         image_size=tuple(meta["image_size"]), correlated=meta["correlated"])
    data = _cuda(data)
    data["num_keypoints0"] = torch.tensor(meta["len0"]).cuda()
    data["num_keypoints1"] = torch.tensor(meta["len1"]).cuda()
    pred = _model(conf, sd)(data)
    M = N = meta["npad"]
    sr, sc = meta["sample_stride"]
    for p in range(meta["batch"]):
        a, c = meta["len0"][p], meta["len1"][p]
        la = pred["log_assignment"][p]
        blk = torch.cat([torch.cat([la[:a, :c], la[:a, N:N + 1]], 1), torch.cat([la[M:M + 1, :c], la[M:M + 1, N:N + 1]], 1)], 0)
        blk = blk.cpu().numpy()  # the reference's (a+1, c+1) matrix of this pair
        err = np.abs(blk[::sr, ::sc] - g[f"log_assignment_sample_{p}"][0]).max()
        err = max(err, np.abs(blk[-3:, :] - g[f"log_assignment_last_{p}"][0]).max())
        rs = np.abs(blk.astype(np.float64).sum(1) - g[f"log_assignment_rowsum_{p}"][0]).max() / (c + 1)
        m0 = pred["matches0"][p, :a].cpu().numpy()
        m1 = pred["matches1"][p, :c].cpu().numpy()
        agree = min((m0 == g[f"matches0_{p}"][0]).mean(), (m1 == g[f"matches1_{p}"][0]).mean())
        n_match = int((g[f"matches0_{p}"][0] >= 0).sum())
        _record("test_forward_varlen_4096_vs_golden", precision=precision, pair=p, la_err=float(err), mean_row_err=float(rs),
                match_agreement=float(agree), reference_matches=n_match)
        assert n_match > 1000  # a configuration without matches would prove nothing about the indices
        assert err < tol and rs < tol
        assert agree > 0.99
        assert (pred["matches0"][p, a:] == -1).all() and (pred["matches1"][p, c:] == -1).all()
        ms = np.abs(pred["matching_scores0"][p, :a].cpu().numpy() - g[f"matching_scores0_{p}"][0]).max()
        assert ms < tol
