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


class PrunedAssignments:
    """log_assignment of a batched adaptive forward: pair b has its own kept counts, so entry b is the reference's
    (1, M'_b + 1, N'_b + 1) matrix over the kept points (lightglue.py:550).  Entries are cut
    out of the padded per-exit-group score tensors on access (len() == batch size)."""

    def __init__(self, n):
        self._where = [None] * n

    def _set(self, b, scores, j, a, c):
        self._where[b] = (scores, j, a, c)

    def __len__(self):
        return len(self._where)

    def __getitem__(self, b):
        scores, j, a, c = self._where[b]
        M, N = scores.shape[1] - 1, scores.shape[2] - 1
        dev = scores.device
        rows = torch.cat([torch.arange(a, device=dev), torch.tensor([M], device=dev)])
        cols = torch.cat([torch.arange(c, device=dev), torch.tensor([N], device=dev)])
        return scores[j].index_select(0, rows).index_select(1, cols)[None]


def forward_adaptive_batched(model, eng, cs0, cs1, len0=None, len1=None):
    """Early exit + token pruning for a BATCH of pairs (the reference asserts b == 1, lightglue.py:528,533; SURVEY.md
    8(f) N2 / config 3).  Every pair follows exactly the reference's per-pair rules -- same thresholds, same stop test
    on its own kept tokens over its own original m + n (lightglue.py:595-606), same order-preserving compaction
    (:536-547) -- while the kernels run on the whole active set: pruned pairs are compacted to the front of their rows
    and carried as valid lengths (the attention / assignment kernels mask by length), pairs that exit are assigned with
    the head of their exit layer and leave the batch, and the survivors are re-tiled to the longest kept length.
    One small device->host read per layer (kept / unconfident counts), like the reference's per-layer Python `if`; the
    compaction for the next layer is queued before that read, so a layer where nobody exits costs the GPU no wait."""
    conf = model.conf
    B, m, n = eng.B, eng.M, eng.N
    dev = eng.dev
    L = conf.n_layers
    do_early_stop = conf.depth_confidence > 0
    do_point_pruning = conf.width_confidence > 0
    min_match = float(np.float32(1 - conf.width_confidence))
    th = conf.filter_threshold

    ids = torch.arange(B, device=dev)                       # original pair id of every active row
    l0h = [m] * B if len0 is None else [int(v) for v in len0.tolist()]
    l1h = [n] * B if len1 is None else [int(v) for v in len1.tolist()]
    denom = [a + c for a, c in zip(l0h, l1h)]               # check_if_stop: ORIGINAL point count of the pair
    l0, l1 = len0, len1
    ind0 = ind1 = pr0 = pr1 = None
    if do_point_pruning:
        ind0 = torch.arange(m, device=dev)[None].repeat(B, 1)
        ind1 = torch.arange(n, device=dev)[None].repeat(B, 1)
        pr0 = torch.zeros(B, m, device=dev, dtype=torch.int64)
        pr1 = torch.zeros(B, n, device=dev, dtype=torch.int64)
        ar0, ar1 = torch.arange(m, device=dev)[None], torch.arange(n, device=dev)[None]
        pr0 += 1 if l0 is None else (ar0 < l0[:, None]).long()  # prune counters start at 1 on real keypoints
        pr1 += 1 if l1 is None else (ar1 < l1[:, None]).long()
    out_m0 = torch.full((B, m + 1), -1, device=dev, dtype=torch.int64)
    out_m1 = torch.full((B, n + 1), -1, device=dev, dtype=torch.int64)
    out_s0 = torch.zeros((B, m + 1), device=dev)
    out_s1 = torch.zeros((B, n + 1), device=dev)
    out_p0 = torch.zeros(B, m, device=dev, dtype=torch.int64) if do_point_pruning else None
    out_p1 = torch.zeros(B, n, device=dev, dtype=torch.int64) if do_point_pruning else None
    exit_layer = torch.full((B,), L - 1, dtype=torch.int64)
    assignments = PrunedAssignments(B)

    def finalize(layer, e, sel, d0, d1, l0e, l1e, i0e, i1e, ide, k0, k1):
        """Assignment + filter of pairs with the head of `layer`, scattered back to the original keypoint indices of
        their pairs (lightglue.py:549-562).  `e`: an engine whose tiles hold exactly these rows, or None (then the rows
        `d0, d1` -- after `sel` -- are tiled into a fresh one)."""
        if sel is not None:
            pick = lambda t: None if t is None else t.index_select(0, sel)
            d0, d1, l0e, l1e, i0e, i1e, ide = map(pick, (d0, d1, l0e, l1e, i0e, i1e, ide))
        if e is None or sel is not None:
            e = model._engine(d0.shape[0], d0.shape[1], d1.shape[1], dev)
            e.load_rows(d0.contiguous(), d1.contiguous())
        scores, (f0, f1, fs0, fs1) = model._assign(layer, e, d0, d1, l0e, l1e, th=th)
        rows = ide[:, None]
        if do_point_pruning:
            mc, nc = d0.shape[1], d1.shape[1]
            t0, t1 = i0e, i1e
            if l0e is not None:  # padding rows land in the spare column
                t0 = torch.where(torch.arange(mc, device=dev)[None] < l0e[:, None], i0e, m)
                t1 = torch.where(torch.arange(nc, device=dev)[None] < l1e[:, None], i1e, n)
            out_m0[rows, t0] = torch.where(f0 == -1, -1, i1e.gather(1, f0.clamp(min=0)))
            out_m1[rows, t1] = torch.where(f1 == -1, -1, i0e.gather(1, f1.clamp(min=0)))
            out_s0[rows, t0] = fs0
            out_s1[rows, t1] = fs1
        else:
            out_m0[ide, : f0.shape[1]] = f0
            out_m1[ide, : f1.shape[1]] = f1
            out_s0[ide, : f0.shape[1]] = fs0
            out_s1[ide, : f1.shape[1]] = fs1
        for j, b in enumerate(ide.tolist()):
            assignments._set(b, scores, j, k0[j], k1[j])
            exit_layer[b] = layer

    counts_host = torch.empty(B, 4, dtype=torch.int32).pin_memory()
    ready = torch.cuda.Event()
    for i in range(L):
        eng.layer(i, l0, l1)
        d0, d1 = eng.rows()
        if i == L - 1:
            finalize(i, eng, None, d0, d1, l0, l1, ind0, ind1, ids, l0h, l1h)
            if do_point_pruning:
                out_p0[ids], out_p1[ids] = pr0, pr1
            break
        thr = float(model.confidence_thresholds[i])
        head = model.log_assignment[i].matchability
        tok = model.token_confidence[i].token[0] if do_early_stop else None
        wt = None if tok is None else tok.weight.reshape(-1)
        bt = None if tok is None else tok.bias
        conf0, match0 = C.token_scores(d0, wt, bt, head.weight.reshape(-1), head.bias, l0)
        conf1, match1 = C.token_scores(d1, wt, bt, head.weight.reshape(-1), head.bias, l1)
        keep0, cnt0 = C.prune_plan(conf0, match0, thr, min_match, l0)
        keep1, cnt1 = C.prune_plan(conf1, match1, thr, min_match, l1)
        nact = d0.shape[0]
        counts_host[:nact].copy_(torch.cat([cnt0, cnt1], 1), non_blocking=True)  # (kept0, unconf0, kept1, unconf1)
        ready.record()
        g0 = g1 = None
        if do_point_pruning:
            # Speculation, queued BEFORE the host waits for the counts: order-preserving compaction of every pair
            # + prune[ind] += 1 on the kept points (lightglue.py:536-547), into the same tile geometry with the new
            # lengths on the device.  If nobody exits and no tile pair is freed (the common layer), the next layer
            # starts right behind it -- the GPU never waits for the host.
            pr0_prev, pr1_prev = pr0.clone(), pr1.clone()
            g0 = C.prune_gather(d0, cs0[0], cs0[1], ind0, keep0, cnt0, pr0, zero=True)
            g1 = C.prune_gather(d1, cs1[0], cs1[1], ind1, keep1, cnt1, pr1, zero=True)
            l0n, l1n = cnt0[:, 0].contiguous(), cnt1[:, 0].contiguous()
            eng.load_rows(g0[0], g1[0], (g0[1], g0[2]), (g1[1], g1[2]))
        ready.synchronize()  # the one host read of this layer
        counts = counts_host[:nact].tolist()
        stop = [False] * nact
        if do_early_stop:  # check_if_stop (lightglue.py:595-606) per pair, fp32 arithmetic
            idl = ids.tolist()
            for j, (k0, u0, k1, u1) in enumerate(counts):
                ratio = np.float32(1.0) - np.float32(u0 + u1) / np.float32(denom[idl[j]])
                stop[j] = bool(ratio > np.float32(conf.depth_confidence))
        E = [j for j in range(nact) if stop[j]]
        R = [j for j in range(nact) if not stop[j]]
        k0 = [counts[j][0] for j in R]
        k1 = [counts[j][2] for j in R]
        if do_point_pruning and R and (min(k0) == 0 or min(k1) == 0):
            raise RuntimeError("every keypoint of one image was pruned (the reference's max() would raise too)")
        if not E and (not do_point_pruning or (C.tiles_pad(max(k0)) == eng.tp0 and C.tiles_pad(max(k1)) == eng.tp1)):
            if do_point_pruning:  # the speculated state is the state
                cs0, cs1, ind0, ind1 = (g0[1], g0[2]), (g1[1], g1[2]), g0[3], g1[3]
                l0, l1, l0h, l1h = l0n, l1n, k0, k1
            continue
        if E:  # exits are assigned from the rows of THIS layer, before its pruning (the reference breaks first)
            sel = None if not R else torch.tensor(E, device=dev)
            finalize(i, None, sel, d0, d1, l0, l1, ind0, ind1, ids, [l0h[j] for j in E], [l1h[j] for j in E])
            if do_point_pruning:
                ide = ids if sel is None else ids.index_select(0, sel)
                out_p0[ide] = pr0_prev if sel is None else pr0_prev.index_select(0, sel)
                out_p1[ide] = pr1_prev if sel is None else pr1_prev.index_select(0, sel)
        if not R:
            break
        # survivors, in order: re-tiled to the longest kept length
        sel = torch.tensor(R, device=dev) if E else None
        pick = lambda t: t if (t is None or sel is None) else t.index_select(0, sel)
        ids = pick(ids)
        if do_point_pruning:
            mk, nk = max(k0), max(k1)
            cut = lambda t, k: pick(t)[:, :k].contiguous()
            d0, cs0, ind0 = cut(g0[0], mk), (cut(g0[1], mk), cut(g0[2], mk)), cut(g0[3], mk)
            d1, cs1, ind1 = cut(g1[0], nk), (cut(g1[1], nk), cut(g1[2], nk)), cut(g1[3], nk)
            pr0, pr1, l0, l1, l0h, l1h = pick(pr0), pick(pr1), pick(l0n), pick(l1n), k0, k1
        else:
            d0, d1, l0, l1 = pick(d0).contiguous(), pick(d1).contiguous(), pick(l0), pick(l1)
            cs0, cs1 = (pick(cs0[0]), pick(cs0[1])), (pick(cs1[0]), pick(cs1[1]))
            l0h, l1h = [l0h[j] for j in R], [l1h[j] for j in R]
        eng = model._engine(len(R), d0.shape[1], d1.shape[1], dev)
        eng.load_rows(d0, d1, cs0, cs1)

    pred = {
        "matches0": out_m0[:, :m].contiguous(), "matches1": out_m1[:, :n].contiguous(),
        "matching_scores0": out_s0[:, :m].contiguous(), "matching_scores1": out_s1[:, :n].contiguous(),
        "log_assignment": assignments, "exit_layer": exit_layer,
    }
    if do_point_pruning:
        pred["prune0"], pred["prune1"] = out_p0, out_p1
    else:
        pred["prune0"] = torch.ones_like(pred["matching_scores0"]) * L
        pred["prune1"] = torch.ones_like(pred["matching_scores1"]) * L
    # descriptors of the last layer exist only for pairs that ran all layers: left out of the batched contract
    pred["ref_descriptors0"] = out_s0.new_zeros((B, 0, m, conf.descriptor_dim))
    pred["ref_descriptors1"] = out_s1.new_zeros((B, 0, n, conf.descriptor_dim))
    return pred
