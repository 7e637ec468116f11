"""Adaptive depth / width forward (early exit + token pruning), reference
gluefactory/models/matchers/lightglue.py:502-565 in eval mode.

The reference only supports one pair per call here (`assert b == 1`, lightglue.py:528,533) and
synchronises with the host once per layer (a Python `if` on a tensor, and `torch.where`); this
path keeps both properties: one 16-byte device->host read per layer carries the kept counts and
the confidence counts of both images.  Intended semantics where the fork is broken
(SURVEY.md 8(a) A12): thresholds from confidence_threshold(i); early exit returns a result.
"""
import numpy as np
import torch

from . import _cabi as C


def forward_adaptive(model, eng, cs0, cs1):
    """eng: TileEngine already loaded with the (projected) descriptors of ONE pair."""
    conf = model.conf
    b, m, n = eng.B, eng.M, eng.N
    assert b == 1, "adaptive depth/width runs one pair per call, like the reference (lightglue.py:528,533)"
    dev = eng.dev
    L = conf.n_layers
    do_early_stop = conf.depth_confidence > 0
    do_point_pruning = conf.width_confidence > 0
    if do_point_pruning:
        ind0 = torch.arange(m, device=dev)[None].contiguous()
        ind1 = torch.arange(n, device=dev)[None].contiguous()
        prune0, prune1 = torch.ones_like(ind0), torch.ones_like(ind1)
    min_match = float(np.float32(1 - conf.width_confidence))
    last = L - 1
    ref_desc = None
    for i in range(L):
        eng.layer(i)
        desc0, desc1 = eng.rows()
        if i == L - 1:
            ref_desc = (desc0, desc1)
            break
        thr = float(model.confidence_thresholds[i])
        head = model.log_assignment[i].matchability
        tok = model.token_confidence[i].token[0] if do_early_stop else None
        wt = None if tok is None else tok.weight.reshape(-1)
        bt = None if tok is None else tok.bias
        conf0, match0 = C.token_scores(desc0, wt, bt, head.weight.reshape(-1), head.bias)
        conf1, match1 = C.token_scores(desc1, wt, bt, head.weight.reshape(-1), head.bias)
        keep0, cnt0 = C.prune_plan(conf0, match0, thr, min_match)
        keep1, cnt1 = C.prune_plan(conf1, match1, thr, min_match)
        k0, u0, k1, u1 = torch.cat([cnt0, cnt1], 1)[0].tolist()  # the one host sync of this layer
        if do_early_stop:
            # check_if_stop (lightglue.py:595-606): fp32 arithmetic, denominator = ORIGINAL m + n
            ratio = np.float32(1.0) - np.float32(u0 + u1) / np.float32(m + n)
            if ratio > np.float32(conf.depth_confidence):
                last = i
                break
        if do_point_pruning and (k0 < eng.M or k1 < eng.N):
            if k0 == 0 or k1 == 0:
                raise RuntimeError("every keypoint of one image was pruned (the reference's max() would raise too)")
            desc0, c0, s0, ind0 = C.prune_gather(desc0, cs0[0], cs0[1], ind0, keep0, cnt0, prune0)
            desc1, c1, s1, ind1 = C.prune_gather(desc1, cs1[0], cs1[1], ind1, keep1, cnt1, prune1)
            desc0, cs0, ind0 = desc0[:, :k0].contiguous(), (c0[:, :k0].contiguous(), s0[:, :k0].contiguous()), ind0[:, :k0].contiguous()
            desc1, cs1, ind1 = desc1[:, :k1].contiguous(), (c1[:, :k1].contiguous(), s1[:, :k1].contiguous()), ind1[:, :k1].contiguous()
            # re-tile the compacted tokens (the index_select of lightglue.py:537-547)
            eng = model._engine(1, k0, k1, dev)
            eng.load_rows(desc0, desc1, cs0, cs1)
        elif do_point_pruning:  # nothing dropped: only the prune counters advance (lightglue.py:540,547)
            prune0[:, ind0[0]] += 1
            prune1[:, ind1[0]] += 1
    scores, (m0, m1, ms0, ms1) = model._assign(last, eng, desc0, desc1, None, None, th=conf.filter_threshold)
    if do_point_pruning:  # scatter back through ind0 / ind1 (lightglue.py:553-562)
        m0_ = torch.full((b, m), -1, device=dev, dtype=m0.dtype)
        m1_ = torch.full((b, n), -1, device=dev, dtype=m1.dtype)
        m0_[:, ind0[0]] = torch.where(m0 == -1, -1, ind1.gather(1, m0.clamp(min=0)))
        m1_[:, ind1[0]] = torch.where(m1 == -1, -1, ind0.gather(1, m1.clamp(min=0)))
        ms0_ = torch.zeros((b, m), device=dev)
        ms1_ = torch.zeros((b, n), device=dev)
        ms0_[:, ind0[0]] = ms0
        ms1_[:, ind1[0]] = ms1
        m0, m1, ms0, ms1 = m0_, m1_, ms0_, ms1_
    else:
        prune0 = torch.ones_like(ms0) * L
        prune1 = torch.ones_like(ms1) * L
    pred = {
        "matches0": m0, "matches1": m1, "matching_scores0": ms0, "matching_scores1": ms1,
        "log_assignment": scores, "prune0": prune0, "prune1": prune1,
        "exit_layer": last,
    }
    if ref_desc is not None:
        pred["ref_descriptors0"], pred["ref_descriptors1"] = ref_desc[0][:, None], ref_desc[1][:, None]
    else:  # early exit: the reference's torch.stack([]) has nothing to return (A12 ii)
        pred["ref_descriptors0"] = desc0.new_zeros((b, 0, desc0.shape[1], desc0.shape[2]))
        pred["ref_descriptors1"] = desc1.new_zeros((b, 0, desc1.shape[1], desc1.shape[2]))
    return pred
