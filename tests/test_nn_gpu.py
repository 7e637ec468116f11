"""GPU parity of the NearestNeighborMatcher drop-in against the reference's golden runs and the
oracle (oracle/nn_oracle.py)."""
import numpy as np
import pytest
import torch

from conftest import load_golden, pkg
from oracle import nn_oracle

pytestmark = pytest.mark.gpu
W = pkg("weights")
NN_CASES = ["nn_mutual_b2_300x257", "nn_ratio_b1_512x400", "nn_dist_nomutual_b1_130x200", "nn_all_b1_1024"]


def _cuda(data):
    return {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda()) for k, v in data.items()}


@pytest.mark.parametrize("name", NN_CASES)
def test_nn_matcher_vs_golden(name):
    meta, g = load_golden(name)
    data = W.make_inputs(meta["batch"], meta["m"], meta["n"], seed=meta["seed"], correlated=meta["correlated"])
    model = pkg("nn_matcher").NearestNeighborMatcher(meta["conf"])
    pred = model(_cuda(data))
    # indices and 0/1 scores: exact, except where a threshold test sits within the similarity's rounding (TF32x3
    # on the GPU, 2e-6 abs, vs the CPU's fp32 einsum)
    for k in ("matches0", "matches1", "matching_scores0", "matching_scores1"):
        got = pred[k].cpu().numpy()
        assert got.dtype == g[k].dtype and got.shape == g[k].shape, k
        agree = (got == g[k]).mean()
        print(f"{name} {k}: agreement {agree:.5f}")
        assert agree >= 0.998, (k, agree)
    assert np.abs(pred["similarity"][:, ::7, ::5].cpu().numpy() - g["similarity_sample"]).max() < 5e-6
    assert np.abs(pred["log_assignment"][:, ::7, ::5].cpu().numpy() - g["log_assignment_sample"]).max() < 3e-5
    la = pred["log_assignment"]
    assert (la[:, -1, :] == 0).all() and (la[:, :, -1] == 0).all()


def test_nn_match_on_reference_similarity_bit_exact():
    """find_nn + mutual_check fed the reference's own similarity tensor: indices bit-exact (incl. ties -> lowest index)."""
    C = pkg("_cabi")
    g = torch.Generator().manual_seed(3)
    sim = torch.randn(2, 333, 190, generator=g).clamp(-1, 1) * 0.9
    sim[0, 5] = sim[0, 4]          # equal rows
    sim[1, :, 7] = sim[1, :, 3]    # equal columns -> ties in every row
    for conf in ({}, {"ratio_thresh": 0.97}, {"distance_thresh": 1.0, "mutual_check": False},
                 {"ratio_thresh": 0.99, "distance_thresh": 1.2}):
        m0 = nn_oracle.find_nn(sim, conf.get("ratio_thresh"), conf.get("distance_thresh"))
        m1 = nn_oracle.find_nn(sim.transpose(1, 2), conf.get("ratio_thresh"), conf.get("distance_thresh"))
        if conf.get("mutual_check", True):
            m0, m1 = nn_oracle.mutual_check(m0, m1)
        got = C.nn_match(sim.cuda().contiguous(), conf.get("ratio_thresh"), conf.get("distance_thresh"),
                         conf.get("mutual_check", True))
        assert torch.equal(got[0].cpu(), m0) and torch.equal(got[1].cpu(), m1), conf
        assert torch.equal(got[2].cpu(), (m0 > -1).float())


def test_nn_registry_alias():
    import importlib

    mod = importlib.import_module("b200glue.nn_matcher")
    assert mod.__main_model__ is pkg("nn_matcher").NearestNeighborMatcher
    with pytest.raises(RuntimeError):
        mod.__main_model__({})({"descriptors0": torch.zeros(1, 4, 256), "descriptors1": torch.zeros(1, 4, 256)})


def test_nn_matcher_large_batch_equals_one_pair_calls():
    """16 x 2048 x 2048 through the NN matcher (similarity tiles, double log-softmax, top-2 passes) against one pair
    per call: the persistent kernels run several work items per CTA only at this size."""
    conf = {"ratio_thresh": 0.95, "distance_thresh": 1.2, "mutual_check": True}
    model = pkg("nn_matcher").NearestNeighborMatcher(conf)
    g = torch.Generator().manual_seed(3)
    d0 = torch.nn.functional.normalize(torch.randn(16, 2048, 256, generator=g), dim=-1).cuda()
    d1 = torch.nn.functional.normalize(0.8 * d0.cpu()[:, torch.randperm(2048, generator=g)] +
                                       0.02 * torch.randn(16, 2048, 256, generator=g), dim=-1).cuda()
    pred = model({"descriptors0": d0, "descriptors1": d1})
    assert (pred["matches0"] >= 0).float().mean() > 0.5
    for b in (0, 9, 15):
        one = model({"descriptors0": d0[b:b + 1], "descriptors1": d1[b:b + 1]})
        assert torch.equal(one["matches0"][0], pred["matches0"][b]) and torch.equal(one["matches1"][0], pred["matches1"][b])
        assert (one["log_assignment"][0] - pred["log_assignment"][b]).abs().max() < 1e-5


def test_nn_matcher_128d_descriptors_vs_oracle():
    """The reference's shipped NN configs (disk+NN, aliked+NN) use 128-d descriptors: any width up to 256 is taken
    (zero-padded to the kernel's 256 columns); wider ones are refused with a clear error."""
    conf = {"ratio_thresh": 0.9, "mutual_check": True}
    data = W.make_inputs(2, 300, 257, d_in=128, seed=21, correlated=0.7)
    ref = nn_oracle.forward(conf, data)
    pred = pkg("nn_matcher").NearestNeighborMatcher(conf)(_cuda(data))
    assert pred["similarity"].shape == ref["similarity"].shape
    assert (pred["similarity"].cpu() - ref["similarity"]).abs().max() < 5e-6
    assert (pred["log_assignment"].cpu() - ref["log_assignment"]).abs().max() < 3e-5
    for k in ("matches0", "matches1"):
        assert (pred[k].cpu() == ref[k]).float().mean() >= 0.998, k
    wide = {"descriptors0": torch.zeros(1, 8, 320).cuda(), "descriptors1": torch.zeros(1, 8, 320).cuda()}
    with pytest.raises(ValueError, match="width"):
        pkg("nn_matcher").NearestNeighborMatcher(conf)(wide)
