"""Adaptive depth / width forward (early exit + token pruning), reference
gluefactory/models/matchers/lightglue.py:502-565 in eval mode.

The reference only supports one pair per call here (`assert b == 1`, lightglue.py:528,533) and
synchronises with the host once per layer (a Python `if` on a tensor, and `torch.where`).  This
module runs a BATCH of pairs with the reference's per-pair rules; one pair per call is the batch
of one.  Per layer: one fused launch on the engine's tiles (lg_score_prune_tiles) and one
16-byte-per-pair device->host read (kept counts and confidence counts of both images).  Intended
semantics where the fork is broken: thresholds from confidence_threshold(i);
early exit returns a result.
"""
import numpy as np
import torch

from . import _cabi as C


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


def _sub_engine(model, eng, sel, x_src, rot_src, mk, nk):
    """A fresh engine for the images `sel` (None = all) of `eng`, cut to mk / nk rows: whole row tiles are copied
    (torch copies of tile blocks -- the rows inside were already compacted / zero-filled by lg_score_prune_tiles)."""
    nact = eng.B
    nb = nact if sel is None else int(sel.numel())
    e = model._engine(nb, mk, nk, eng.dev)
    pick = (lambda v: v) if sel is None else (lambda v: v.index_select(0, sel))
    for src0, tps, dst0, tpd in ((0, eng.tp0, 0, e.tp0), (eng.rt0, eng.tp1, e.rt0, e.tp1)):
        for src, dst in ((x_src, e.x), (rot_src, e.rot)):
            blk = src[src0: src0 + nact * tps].view(nact, tps, -1)
            dst[dst0: dst0 + nb * tpd].view(nb, tpd, -1).copy_(pick(blk)[:, :tpd])
    return e


def forward_adaptive_batched(model, eng, cs0, cs1, len0=None, len1=None):
    """Early exit + token pruning for a BATCH of pairs (the reference asserts b == 1, lightglue.py:528,533; SURVEY.md
    8(f) N2 / config 3).  Every pair follows exactly the reference's per-pair rules -- same thresholds, same stop test
    on its own kept tokens over its own original m + n (lightglue.py:595-606), same order-preserving compaction
    (:536-547) -- while the kernels run on the whole active set.  The adaptive step of a layer is ONE launch on the
    engine's split tiles (lg_score_prune_tiles: scores, keep mask, scan, compaction into the engine's second buffer,
    new lengths on the device) + one 16*B-byte device->host read (kept / unconfident counts), like the reference's
    per-layer Python `if`.  Pruned pairs are carried as valid lengths (the attention / assignment kernels mask by
    length); pairs that exit are assigned with the head of their exit layer from the tiles as they were BEFORE this
    layer's pruning (the reference breaks first) and leave the batch; the survivors move to a smaller engine only when
    a pair left or a whole tile pair became free."""
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
    ind0 = ind1 = pr0 = pr1 = None                          # ind None = identity (nothing pruned yet)
    if do_point_pruning:
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
    ref_desc = None

    def finalize(layer, e, l0e, l1e, i0e, i1e, ide, k0, k1):
        """Assignment + filter of the pairs held by engine `e` with the head of `layer`, scattered back to the original
        keypoint indices of their pairs (lightglue.py:549-562)."""
        d0, d1 = e.rows()
        scores, (f0, f1, fs0, fs1) = model._assign(layer, e, d0, d1, l0e, l1e, th=th)
        rows = ide[:, None]
        mc, nc = e.M, e.N
        if do_point_pruning:
            if i0e is None:
                i0e = torch.arange(mc, device=dev)[None].expand(e.B, mc)
                i1e = torch.arange(nc, device=dev)[None].expand(e.B, nc)
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
        return d0, d1

    sets = {}

    def spare(nact, mc, nc):
        """Output buffers of the next compaction: two sets per shape, used in turn (the set written last holds the
        current state, the other one is free again -- the kernel overwrites every entry it hands out)."""
        key = (nact, mc, nc)
        if key not in sets:
            mk = lambda: (torch.empty(nact, mc, device=dev, dtype=torch.int64), torch.empty(nact, nc, device=dev, dtype=torch.int64),
                          torch.empty(nact, m, device=dev, dtype=torch.int64), torch.empty(nact, n, device=dev, dtype=torch.int64),
                          torch.empty(nact, device=dev, dtype=torch.int32), torch.empty(nact, device=dev, dtype=torch.int32))
            sets[key] = [mk(), mk(), 0]
        entry = sets[key]
        entry[2] ^= 1
        return entry[entry[2]]

    counts_host = model.__dict__.get("_lg_counts_host")  # pinned once per model (pinning costs ~1 ms)
    if counts_host is None or counts_host.shape[0] < B:
        counts_host = model.__dict__["_lg_counts_host"] = torch.empty(max(B, 64), 4, dtype=torch.int32).pin_memory()
    ready = torch.cuda.Event()
    for i in range(L):
        eng.layer(i, l0, l1)
        if i == L - 1:
            d0, d1 = finalize(i, eng, l0, l1, ind0, ind1, ids, l0h, l1h)
            if eng.B == B and not do_point_pruning:  # nobody left, nothing pruned: the reference's ref_descriptors
                ref_desc = (d0, d1)
            elif B == 1:  # one pair that ran all layers: the descriptors of its kept points (lightglue.py:572-573)
                ref_desc = (d0[:, : l0h[0]], d1[:, : l1h[0]])
            if do_point_pruning:
                out_p0[ids], out_p1[ids] = pr0, pr1
            break
        thr = float(model.confidence_thresholds[i])
        head = model.log_assignment[i].matchability
        tok = model.token_confidence[i].token[0] if do_early_stop else None
        wt = None if tok is None else tok.weight.reshape(-1)
        bt = None if tok is None else tok.bias
        nact = eng.B
        counts = torch.empty(nact, 4, device=dev, dtype=torch.int32)
        if do_point_pruning:
            # Speculation, queued BEFORE the host waits for the counts: the compacted state of every pair goes to the
            # engine's second buffers; if nobody exits and no tile pair is freed (the common layer) the buffers are
            # swapped and the next layer starts right behind -- the GPU never waits for the host.
            ind0n, ind1n, pr0n, pr1n, l0n, l1n = spare(nact, eng.M, eng.N)
            C.score_prune_tiles(eng.x, eng.rot, eng.x_alt, eng.rot_alt(), (ind0, ind1), (ind0n, ind1n), (pr0, pr1),
                                (pr0n, pr1n), (l0, l1), (l0n, l1n), counts, wt, bt, head.weight.reshape(-1), head.bias,
                                thr, min_match, True, nact, eng.M, eng.N, eng.tp0, eng.tp1, m, n)
        else:
            C.score_prune_tiles(eng.x, None, None, None, None, None, None, None, (l0, l1), None, counts, wt, bt,
                                head.weight.reshape(-1), head.bias, thr, min_match, False, nact, eng.M, eng.N, eng.tp0,
                                eng.tp1, m, n)
        counts_host[:nact].copy_(counts, non_blocking=True)  # (kept0, unconf0, kept1, unconf1)
        ready.record()
        ready.synchronize()  # the one host read of this layer
        cnt = counts_host[:nact].tolist()
        stop = [False] * nact
        if do_early_stop:  # check_if_stop (lightglue.py:595-606) per pair, fp32 arithmetic
            idl = ids.tolist()
            for j, (k0, u0, k1, u1) in enumerate(cnt):
                ratio = np.float32(1.0) - np.float32(u0 + u1) / np.float32(denom[idl[j]])
                stop[j] = bool(ratio > np.float32(conf.depth_confidence))
        E = [j for j in range(nact) if stop[j]]
        R = [j for j in range(nact) if not stop[j]]
        k0 = [cnt[j][0] for j in R]
        k1 = [cnt[j][2] for j in R]
        if do_point_pruning and R and (min(k0) == 0 or min(k1) == 0):
            raise RuntimeError("every keypoint of one image was pruned (the reference's max() would raise too)")
        same_tiles = (not R) or (not do_point_pruning) or (C.tiles_pad(max(k0)) == eng.tp0 and C.tiles_pad(max(k1)) == eng.tp1)
        if not E and same_tiles:
            if do_point_pruning:  # the speculated state is the state
                eng.swap_pruned()
                ind0, ind1, pr0, pr1, l0, l1, l0h, l1h = ind0n, ind1n, pr0n, pr1n, l0n, l1n, k0, k1
            continue
        if E:  # exits are assigned from the tiles of THIS layer, before its pruning (the reference breaks first)
            sel = None if not R else torch.tensor(E, device=dev)
            pick = lambda t: t if (t is None or sel is None) else t.index_select(0, sel)
            e_exit = eng if sel is None else _sub_engine(model, eng, sel, eng.x, eng.rot, eng.M, eng.N)
            finalize(i, e_exit, pick(l0), pick(l1), pick(ind0), pick(ind1), pick(ids), [l0h[j] for j in E], [l1h[j] for j in E])
            if do_point_pruning:
                out_p0[pick(ids)], out_p1[pick(ids)] = pick(pr0), pick(pr1)
        if not R:
            break
        # survivors, in order: a smaller engine (fewer pairs and / or fewer row tiles)
        sel = torch.tensor(R, device=dev) if E else None
        pick = lambda t: t if (t is None or sel is None) else t.index_select(0, sel)
        ids = pick(ids)
        if do_point_pruning:
            mk, nk = max(k0), max(k1)
            eng = _sub_engine(model, eng, sel, eng.x_alt, eng.rot_alt(), mk, nk)
            ind0, ind1 = pick(ind0n)[:, :mk].contiguous(), pick(ind1n)[:, :nk].contiguous()
            pr0, pr1, l0, l1, l0h, l1h = pick(pr0n), pick(pr1n), pick(l0n), pick(l1n), k0, k1
        else:
            eng = _sub_engine(model, eng, sel, eng.x, eng.rot, eng.M, eng.N)
            l0, l1 = pick(l0), pick(l1)
            l0h, l1h = [l0h[j] for j in R], [l1h[j] for j in R]

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
    if ref_desc is not None:  # every pair ran all layers on all its points, as in the reference's full-depth return
        pred["ref_descriptors0"], pred["ref_descriptors1"] = ref_desc[0][:, None], ref_desc[1][:, None]
    else:  # pairs left early or lost points: per-pair descriptors have no common shape (the reference: A12 ii)
        pred["ref_descriptors0"] = out_s0.new_zeros((B, 0, m, conf.descriptor_dim))
        pred["ref_descriptors1"] = out_s1.new_zeros((B, 0, n, conf.descriptor_dim))
    return pred


def forward_adaptive_one(model, eng, cs0, cs1):
    """One pair per call (the reference's only adaptive mode) through the batched path: same kernels, the reference's
    output contract -- log_assignment a (1, M'+1, N'+1) tensor, exit_layer an int."""
    pred = forward_adaptive_batched(model, eng, cs0, cs1)
    pred["log_assignment"] = pred["log_assignment"][0]
    pred["exit_layer"] = int(pred["exit_layer"][0])
    return pred
