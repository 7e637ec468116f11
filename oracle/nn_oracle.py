"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

CPU restatement of the reference's NearestNeighborMatcher forward
(gluefactory/models/matchers/nearest_neighbor_matcher.py), pinned bit-exactly against the reference module run in the
build container (oracle/make_golden_nn.py -> tests/golden/nn_*.npz, tests/test_oracle_golden.py)."""
import torch
import torch.nn.functional as F


def find_nn(sim, ratio_thresh, distance_thresh):
    """nearest_neighbor_matcher.py:16-27."""
    k = 2 if ratio_thresh else 1
    sim_nn, ind_nn = sim.topk(k, dim=-1, largest=True)
    dist_nn = 2 * (1 - sim_nn)
    mask = torch.ones(ind_nn.shape[:-1], dtype=torch.bool)
    if ratio_thresh:
        mask = mask & (dist_nn[..., 0] <= (ratio_thresh ** 2) * dist_nn[..., 1])
    if distance_thresh:
        mask = mask & (dist_nn[..., 0] <= distance_thresh ** 2)
    return torch.where(mask, ind_nn[..., 0], torch.full_like(ind_nn[..., 0], -1))


def mutual_check(m0, m1):
    """nearest_neighbor_matcher.py:30-39."""
    inds0, inds1 = torch.arange(m0.shape[-1]), torch.arange(m1.shape[-1])
    loop0 = torch.gather(m1, -1, torch.where(m0 > -1, m0, torch.zeros_like(m0)))
    loop1 = torch.gather(m0, -1, torch.where(m1 > -1, m1, torch.zeros_like(m1)))
    m0_new = torch.where((m0 > -1) & (inds0 == loop0), m0, torch.full_like(m0, -1))
    m1_new = torch.where((m1 > -1) & (inds1 == loop1), m1, torch.full_like(m1, -1))
    return m0_new, m1_new


def forward(conf: dict, data: dict) -> dict:
    """NearestNeighborMatcher._forward, nearest_neighbor_matcher.py:52-72."""
    ratio, dist, mutual = conf.get("ratio_thresh"), conf.get("distance_thresh"), conf.get("mutual_check", True)
    sim = torch.einsum("bnd,bmd->bnm", data["descriptors0"], data["descriptors1"])
    m0 = find_nn(sim, ratio, dist)
    m1 = find_nn(sim.transpose(1, 2), ratio, dist)
    if mutual:
        m0, m1 = mutual_check(m0, m1)
    b, m, n = sim.shape
    la = sim.new_zeros(b, m + 1, n + 1)
    la[:, :-1, :-1] = F.log_softmax(sim, -1) + F.log_softmax(sim, -2)
    return {"matches0": m0, "matches1": m1, "matching_scores0": (m0 > -1).float(), "matching_scores1": (m1 > -1).float(),
            "similarity": sim, "log_assignment": la}
