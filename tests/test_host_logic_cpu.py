"""CPU tests of the host-side logic that needs no GPU: conf merging, weight permutations of the fp32 path, the
per-pair view of a batched adaptive log_assignment, chunking of the fp32 path, error behaviour without CUDA."""
import pytest
import torch

from conftest import pkg

M = pkg("matcher")
A = pkg("adaptive")
E32 = pkg("engine_fp32")
NN = pkg("nn_matcher")


def test_merge_conf_defaults_unknown_keys_and_nesting():
    conf = M.merge_conf(M.LightGlue.default_conf, {"filter_threshold": 0.3, "b200": {"precision": "fp32"}, "extra": {"a": 1}})
    assert conf.filter_threshold == 0.3 and conf.n_layers == 9
    assert conf.b200.precision == "fp32" and conf.b200.cuda_graph is True     # nested defaults survive
    assert conf.extra.a == 1                                                  # unknown keys tolerated, attribute access
    with pytest.raises(ValueError):
        M.LightGlue({"b200": {"precision": "fp8"}})
    with pytest.raises(ValueError):
        M.LightGlue({"num_heads": 8})


def test_state_dict_names_match_reference_layout():
    model = M.LightGlue({"input_dim": 128, "add_scale_ori": True})
    sd = model.state_dict()
    assert sd["posenc.Wr.weight"].shape == (32, 4) and sd["input_proj.weight"].shape == (256, 128)
    assert sd["transformers.8.self_attn.Wqkv.weight"].shape == (768, 256)
    assert sd["transformers.0.cross_attn.to_qk.weight"].shape == (256, 256)
    assert sd["log_assignment.8.matchability.weight"].shape == (1, 256)
    assert "token_confidence.7.token.0.weight" in sd and "token_confidence.8.token.0.weight" not in sd
    legacy = {"self_attn.3.Wqkv.weight": 1, "cross_attn.0.to_v.bias": 2, "posenc.Wr.weight": 3}
    renamed = M.LightGlue.rename_legacy_keys(legacy, 9)
    assert set(renamed) == {"transformers.3.self_attn.Wqkv.weight", "transformers.0.cross_attn.to_v.bias", "posenc.Wr.weight"}


def test_fp32_qkv_permutation_matches_reference_unflatten():
    """lightglue.py:185: qkv.unflatten(-1, (H, -1, 3)).transpose(1, 2) then q, k, v = qkv[..., 0], [..., 1], [..., 2]:
    feature h*192 + d*3 + part.  The fp32 path permutes the Wqkv rows to [q | k | v] blocks with head h at h*64."""
    perm = E32._qkv_perm("cpu")
    y = torch.arange(768.)[None, None]                      # a fake Wqkv output whose value is its feature index
    qkv = y.unflatten(-1, (4, -1, 3)).transpose(1, 2)       # (1, H, 1, 64, 3)
    q, k, v = qkv[..., 0], qkv[..., 1], qkv[..., 2]         # (1, H, 1, 64)
    ours = y[..., perm.long()]                              # what the permuted weight rows produce
    for part, ref in enumerate((q, k, v)):
        for h in range(4):
            assert torch.equal(ours[0, 0, part * 256 + h * 64: part * 256 + (h + 1) * 64], ref[0, h, 0])


def test_fp32_chunking_keeps_scores_under_budget():
    assert E32.max_pairs(2048, 2048) * 4 * 2048 * 2051 <= 1 << 28
    assert E32.max_pairs(8192, 8192) == 1 and E32.max_pairs(128, 100) >= 32


def test_pruned_assignments_view():
    """pair b of a padded score tensor (B, M+1, N+1) with a kept rows / c kept columns -> (1, a+1, c+1): the valid block,
    the dustbin column (index N) and the dustbin row (index M)."""
    B, Mp, Np = 2, 6, 5
    scores = torch.arange(B * (Mp + 1) * (Np + 1), dtype=torch.float32).view(B, Mp + 1, Np + 1)
    pa = A.PrunedAssignments(3)
    pa._set(2, scores, 1, 4, 3)
    got = pa[2]
    assert len(pa) == 3 and got.shape == (1, 5, 4)
    assert torch.equal(got[0, :4, :3], scores[1, :4, :3])
    assert torch.equal(got[0, :4, 3], scores[1, :4, Np]) and torch.equal(got[0, 4, :3], scores[1, Mp, :3])
    assert got[0, 4, 3] == scores[1, Mp, Np]


def test_cpu_inputs_are_rejected_not_computed():
    data = {"keypoints0": torch.zeros(1, 4, 2), "keypoints1": torch.zeros(1, 4, 2),
            "descriptors0": torch.zeros(1, 4, 256), "descriptors1": torch.zeros(1, 4, 256)}
    with pytest.raises(RuntimeError):
        M.LightGlue({})(data)
    with pytest.raises(RuntimeError):
        NN.NearestNeighborMatcher({})(data)
    with pytest.raises(NotImplementedError):
        NN.NearestNeighborMatcher({"loss": "N_pair"})
    assert NN.NearestNeighborMatcher({"ratio_thresh": 0.8}).conf.mutual_check is True


def test_pack_pairs_and_unpack_pred_shapes():
    """pipeline.pack_pairs: zero padding behind the valid points (pad_to_length(mode="zeros"), models/utils/misc.py:19-57),
    counts, per-pair extents when image_size is missing; unpack_pred cuts back to the valid block + dustbins."""
    P = pkg("pipeline")
    W = pkg("weights")
    pairs = [W.make_inputs(1, m, n, seed=3 + i) for i, (m, n) in enumerate([(5, 130), (200, 7)])]
    pairs[0].pop("view0")
    b = P.pack_pairs(pairs)
    assert b["keypoints0"].shape == (2, 256, 2) and b["descriptors1"].shape == (2, 256, 256)
    assert b["num_keypoints0"].tolist() == [5, 200] and b["num_keypoints1"].tolist() == [130, 7]
    assert torch.equal(b["keypoints0"][0, :5], pairs[0]["keypoints0"][0]) and b["keypoints0"][0, 5:].abs().max() == 0
    k = pairs[0]["keypoints0"][0]
    assert torch.equal(b["view0"]["image_size"][0], 1 + k.max(0).values - k.min(0).values)
    assert torch.equal(b["view0"]["image_size"][1], pairs[1]["view0"]["image_size"][0].float())
    assert P.pack_pairs(pairs, length0=384)["keypoints0"].shape[1] == 384
    M, N = 256, 256
    pred = {"matches0": torch.zeros(2, M, dtype=torch.int64), "matches1": torch.zeros(2, N, dtype=torch.int64),
            "matching_scores0": torch.zeros(2, M), "matching_scores1": torch.zeros(2, N),
            "log_assignment": torch.arange(2 * (M + 1) * (N + 1), dtype=torch.float32).reshape(2, M + 1, N + 1)}
    out = P.unpack_pred(pred, b)
    assert out[0]["matches0"].shape == (1, 5) and out[0]["log_assignment"].shape == (1, 6, 131)
    assert out[1]["log_assignment"][0, -1, -1] == pred["log_assignment"][1, M, N]
    assert out[1]["log_assignment"][0, 3, -1] == pred["log_assignment"][1, 3, N]


def test_condition_option_validated():
    with pytest.raises(ValueError):
        M.LightGlue({"b200": {"condition": "per-image"}})
    assert M.LightGlue({}).conf.b200.condition == "padded"


def test_b200glue_pipeline_reexports():
    import importlib
    mod = importlib.import_module("b200glue.pipeline")
    P = pkg("pipeline")
    for name in ("HostPipeline", "pack_pairs", "unpack_pred", "group_by_length", "RESULT_KEYS"):
        assert getattr(mod, name) is getattr(P, name)
    pairs = [{"keypoints0": torch.zeros(1, a, 2), "keypoints1": torch.zeros(1, b, 2)} for a, b in [(5, 9), (100, 3), (7, 7)]]
    groups = P.group_by_length(pairs, 2)
    assert sorted(i for g in groups for i in g) == [0, 1, 2] and groups[0] == [2, 0]
