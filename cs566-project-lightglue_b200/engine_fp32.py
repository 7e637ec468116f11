"""fp32 path of the matcher (`b200.precision: fp32`): the layer stack in IEEE fp32 on fp32 rows.

Same launch structure as the reference's own math branch (gluefactory/models/matchers/lightglue.py:139-156,178-191,
220-249): projections as GEMMs, scores materialised per (pair, head), row softmax, P.v, out projection, FFN with
LayerNorm + GELU and the residual -- every product and sum in fp32 (csrc/lg_fp32.cu), so that log_assignment meets the
north-star's 1e-4 bound against the reference's fp32 run (the tensor-core path: 2e-2 bound, 1.3e-3 measured).  It is
the accuracy mode, not the throughput path (CUDA-core FMA GEMMs); batches are processed in chunks of pairs that keep
the materialised scores under ~1 GB.  Both images share every weight (lightglue.py:268-272): their tokens form ONE
row matrix [image-0 batch | image-1 batch] for all per-token linears.
"""
import torch

from . import _cabi as C

H, DH, D = 4, 64, 256


def _qkv_perm(device):
    """reference Wqkv output row h*192 + d*3 + part (lightglue.py:185) -> row part*256 + h*64 + d."""
    part = torch.arange(3, device=device)[:, None, None]
    h = torch.arange(H, device=device)[None, :, None]
    d = torch.arange(DH, device=device)[None, None, :]
    return (h * 3 * DH + d * 3 + part).reshape(-1)


class Fp32Weights:
    def __init__(self, model):
        dev = next(model.parameters()).device
        f = lambda t: t.detach().float().contiguous()
        perm = _qkv_perm(dev)
        self.layers = []
        for layer in model.transformers:
            sa, ca = layer.self_attn, layer.cross_attn
            s = {"wqkv": f(sa.Wqkv.weight[perm]), "bqkv": f(sa.Wqkv.bias[perm]), "wo": f(sa.out_proj.weight),
                 "bo": f(sa.out_proj.bias)}
            c = {"wqkv": f(torch.cat([ca.to_qk.weight, ca.to_v.weight], 0)),
                 "bqkv": f(torch.cat([ca.to_qk.bias, ca.to_v.bias], 0)), "wo": f(ca.to_out.weight), "bo": f(ca.to_out.bias)}
            for blk, ffn in ((s, sa.ffn), (c, ca.ffn)):
                blk.update(w0=f(ffn[0].weight), b0=f(ffn[0].bias), g=f(ffn[1].weight), be=f(ffn[1].bias),
                           eps=float(ffn[1].eps), w3=f(ffn[3].weight), b3=f(ffn[3].bias))
            self.layers.append((s, c))
        self.final = [(f(h.final_proj.weight), f(h.final_proj.bias), f(h.matchability.weight.reshape(-1)),
                       f(h.matchability.bias)) for h in model.log_assignment]
        self.input_proj = None
        if model.conf.input_dim != D:
            self.input_proj = (f(model.input_proj.weight), f(model.input_proj.bias))


def max_pairs(m: int, n: int, budget_floats: int = 1 << 28) -> int:
    """pairs per chunk so that the materialised scores (B, H, max, pitch) stay under ~1 GB."""
    mx = max(m, n)
    return max(1, budget_floats // (H * mx * (mx + 3)))


class Fp32Engine:
    def __init__(self, w: Fp32Weights, B: int, m: int, n: int, device):
        self.w, self.B, self.m, self.n, self.dev = w, B, m, n, device
        self.T0, self.T = B * m, B * (m + n)
        e = lambda *shape: torch.empty(*shape, device=device, dtype=torch.float32)
        self.x, self.qkv, self.ctx, self.msg, self.h = e(self.T, D), e(self.T, 3 * D), e(self.T, D), e(self.T, D), e(self.T, 2 * D)
        mx = max(m, n)
        self.pitch = (mx + 3) // 4 * 4
        self.s = e(B * H * mx * self.pitch)

    def _linear(self, a, w, b, out, *, n_out, k, lda, ldb, ldc, b_off=0, resid=None):
        C.gemm_f32(a, w, out, self.T, n_out, k, lda, ldb, ldc, bias=b, resid=resid, ldr=ldc, offs=(0, b_off, 0, 0))

    def _attend(self, q_side, k_side, q_col, k_col, v_col, lens):
        """ctx[q side] = softmax(alpha q k^T) v for every pair and head; sides are (row offset, points per image)."""
        (rq, nq), (rk, nk) = q_side, k_side
        B, ld = self.B, 3 * D
        lq, lk = lens
        s = self.s[: B * H * nq * self.pitch].view(B, H, nq, self.pitch)
        C.gemm_f32(self.qkv, self.qkv, s, nq, nk, DH, ld, ld, self.pitch, alpha=DH ** -0.5, batch=B * H, inner=H,
                   strides=(nq * ld, DH, nk * ld, DH, H * nq * self.pitch, nq * self.pitch, 0, 0),
                   offs=(rq * ld + q_col, rk * ld + k_col, 0, 0), m_len=lq)
        C.softmax_rows_f32(s, lq, lk, n=nk)
        C.gemm_f32(s, self.qkv, self.ctx, nq, DH, nk, self.pitch, ld, D, b_kn=True, batch=B * H, inner=H,
                   strides=(H * nq * self.pitch, nq * self.pitch, nk * ld, DH, nq * D, DH, 0, 0),
                   offs=(0, rk * ld + v_col, rq * D, 0), k_len=lk)  # padded query rows: P == 0 -> context 0

    def _ffn(self, blk):
        # out projection, then ffn(cat[x, msg]) as two K=256 GEMMs into h (lightglue.py:190-191,246-248)
        self._linear(self.ctx, blk["wo"], blk["bo"], self.msg, n_out=D, k=D, lda=D, ldb=D, ldc=D)
        self._linear(self.x, blk["w0"], blk["b0"], self.h, n_out=2 * D, k=D, lda=D, ldb=2 * D, ldc=2 * D)
        self._linear(self.msg, blk["w0"], None, self.h, n_out=2 * D, k=D, lda=D, ldb=2 * D, ldc=2 * D, b_off=D, resid=self.h)
        C.ln_gelu_rows_f32(self.h, blk["g"], blk["be"], blk["eps"])
        self._linear(self.h, blk["w3"], blk["b3"], self.x, n_out=D, k=2 * D, lda=2 * D, ldb=2 * D, ldc=D, resid=self.x)

    def load(self, desc0, desc1, cs0, cs1):
        """Descriptors (after input_proj, lightglue.py:478-486) and the rotary tables of both images."""
        B, m, n = self.B, self.m, self.n
        if self.w.input_proj is None or desc0.shape[-1] == D:
            self.x[: self.T0].view(B, m, D).copy_(desc0)
            self.x[self.T0:].view(B, n, D).copy_(desc1)
        else:
            wi, bi = self.w.input_proj
            raw = torch.cat([desc0.reshape(self.T0, -1), desc1.reshape(self.T - self.T0, -1)], 0).contiguous()
            d_in = raw.shape[1]
            C.gemm_f32(raw, wi, self.x, self.T, D, d_in, d_in, d_in, D, bias=bi)
        self.cos = torch.cat([cs0[0].reshape(-1, 32), cs1[0].reshape(-1, 32)], 0).contiguous()
        self.sin = torch.cat([cs0[1].reshape(-1, 32), cs1[1].reshape(-1, 32)], 0).contiguous()

    def layer(self, i, len0=None, len1=None):
        """TransformerLayer i (lightglue.py:258-281) on both images."""
        side0, side1 = (0, self.m), (self.T0, self.n)
        blk_s, blk_c = self.w.layers[i]
        # SelfBlock (lightglue.py:178-191): q | k | v blocks of 256 columns, head h at h*64
        self._linear(self.x, blk_s["wqkv"], blk_s["bqkv"], self.qkv, n_out=3 * D, k=D, lda=D, ldb=D, ldc=3 * D)
        C.rotary_rows_f32(self.qkv.view(1, self.T, 3 * D), self.cos, self.sin, 0, D, H)
        self._attend(side0, side0, 0, D, 2 * D, (len0, len0))
        self._attend(side1, side1, 0, D, 2 * D, (len1, len1))
        self._ffn(blk_s)
        # CrossBlock (lightglue.py:220-249): shared to_qk projection (columns 0..255), to_v (256..511)
        self._linear(self.x, blk_c["wqkv"], blk_c["bqkv"], self.qkv, n_out=2 * D, k=D, lda=D, ldb=D, ldc=3 * D)
        self._attend(side0, side1, 0, 0, D, (len0, len1))
        self._attend(side1, side0, 0, 0, D, (len1, len0))
        self._ffn(blk_c)

    def descriptors(self):
        return self.x[: self.T0].view(self.B, self.m, D), self.x[self.T0:].view(self.B, self.n, D)

    def forward(self, desc0, desc1, cs0, cs1, len0, len1, n_layers):
        self.load(desc0, desc1, cs0, cs1)
        for i in range(n_layers):
            self.layer(i, len0, len1)
        return self.descriptors()

    def assign(self, i, th, len0, len1):
        """MatchAssignment (lightglue.py:306-315) + filter_matches (:321-337): fp32 final_proj, TF32x3 similarity."""
        wf, bf, wm, bm = self.w.final[i]
        md = torch.empty(self.T, D, device=self.dev, dtype=torch.float32)
        self._linear(self.x, wf, bf, md, n_out=D, k=D, lda=D, ldb=D, ldc=D)
        z = C.logit_rows(self.x.view(1, self.T, D), wm, bm)[0]
        B, m, n = self.B, self.m, self.n
        return C.assign_and_filter(md[: self.T0].view(B, m, D), md[self.T0:].view(B, n, D), z[: self.T0].view(B, m),
                                   z[self.T0:].view(B, n), th, len0, len1, prescale=D ** -0.25)


def forward_adaptive_fp32(model, w: Fp32Weights, desc0, desc1, cs0, cs1):
    """Early exit + token pruning (lightglue.py:502-565, eval) of ONE pair in the fp32 mode: the reference's own
    per-layer procedure on fp32 rows -- token confidences and matchability (lg_token_scores), keep mask / counts
    (lg_prune_plan), one 16-byte host read per layer, order-preserving gather of descriptors, rotary rows and original
    indices (lg_prune_gather), then a smaller engine.  Intended semantics of SURVEY.md 8(a) A12 (thresholds from
    confidence_threshold(i); early exit returns a result).  The accuracy mode: not tuned for speed."""
    import numpy as np

    conf = model.conf
    m, n = desc0.shape[1], desc1.shape[1]
    dev = desc0.device
    L = conf.n_layers
    do_early_stop = conf.depth_confidence > 0
    do_point_pruning = conf.width_confidence > 0
    if do_point_pruning:
        ind0 = torch.arange(m, device=dev)[None].contiguous()
        ind1 = torch.arange(n, device=dev)[None].contiguous()
        prune0, prune1 = torch.ones_like(ind0), torch.ones_like(ind1)
    min_match = float(np.float32(1 - conf.width_confidence))
    eng = Fp32Engine(w, 1, m, n, dev)
    eng.load(desc0, desc1, cs0, cs1)
    last, ref_desc = L - 1, None
    for i in range(L):
        eng.layer(i)
        d0, d1 = eng.descriptors()
        if i == L - 1:
            ref_desc = (d0.clone(), d1.clone())
            break
        thr = float(model.confidence_thresholds[i])
        head = model.log_assignment[i].matchability
        tok = model.token_confidence[i].token[0] if do_early_stop else None
        wt = None if tok is None else tok.weight.reshape(-1)
        bt = None if tok is None else tok.bias
        conf0, match0 = C.token_scores(d0, wt, bt, head.weight.reshape(-1), head.bias)
        conf1, match1 = C.token_scores(d1, wt, bt, head.weight.reshape(-1), head.bias)
        keep0, cnt0 = C.prune_plan(conf0, match0, thr, min_match)
        keep1, cnt1 = C.prune_plan(conf1, match1, thr, min_match)
        k0, u0, k1, u1 = torch.cat([cnt0, cnt1], 1)[0].tolist()  # the one host read of this layer
        if do_early_stop:  # check_if_stop (lightglue.py:595-606): fp32 arithmetic, denominator = ORIGINAL m + n
            ratio = np.float32(1.0) - np.float32(u0 + u1) / np.float32(m + n)
            if ratio > np.float32(conf.depth_confidence):
                last = i
                break
        if do_point_pruning and (k0 < eng.m or k1 < eng.n):
            if k0 == 0 or k1 == 0:
                raise RuntimeError("every keypoint of one image was pruned (the reference's max() would raise too)")
            c0, s0 = eng.cos[: eng.m].view(1, eng.m, 32), eng.sin[: eng.m].view(1, eng.m, 32)
            c1, s1 = eng.cos[eng.m:].view(1, eng.n, 32), eng.sin[eng.m:].view(1, eng.n, 32)
            d0, c0, s0, ind0 = C.prune_gather(d0.contiguous(), c0, s0, ind0, keep0, cnt0, prune0)
            d1, c1, s1, ind1 = C.prune_gather(d1.contiguous(), c1, s1, ind1, keep1, cnt1, prune1)
            ind0, ind1 = ind0[:, :k0].contiguous(), ind1[:, :k1].contiguous()
            eng = Fp32Engine(w, 1, k0, k1, dev)
            eng.load(d0[:, :k0], d1[:, :k1], (c0[:, :k0], s0[:, :k0]), (c1[:, :k1], s1[:, :k1]))
        elif do_point_pruning:  # nothing dropped: only the prune counters advance (lightglue.py:540,547)
            prune0[:, ind0[0]] += 1
            prune1[:, ind1[0]] += 1
    scores, (m0, m1, ms0, ms1) = eng.assign(last, conf.filter_threshold, None, None)
    if do_point_pruning:  # scatter back through ind0 / ind1 (lightglue.py:553-562)
        m0_ = torch.full((1, m), -1, device=dev, dtype=m0.dtype)
        m1_ = torch.full((1, n), -1, device=dev, dtype=m1.dtype)
        m0_[:, ind0[0]] = torch.where(m0 == -1, -1, ind1.gather(1, m0.clamp(min=0)))
        m1_[:, ind1[0]] = torch.where(m1 == -1, -1, ind0.gather(1, m1.clamp(min=0)))
        ms0_, ms1_ = torch.zeros((1, m), device=dev), torch.zeros((1, n), device=dev)
        ms0_[:, ind0[0]] = ms0
        ms1_[:, ind1[0]] = ms1
        m0, m1, ms0, ms1 = m0_, m1_, ms0_, ms1_
    else:
        prune0, prune1 = torch.ones_like(ms0) * L, torch.ones_like(ms1) * L
    pred = {"matches0": m0, "matches1": m1, "matching_scores0": ms0, "matching_scores1": ms1, "log_assignment": scores,
            "prune0": prune0, "prune1": prune1, "exit_layer": last}
    if ref_desc is not None:
        pred["ref_descriptors0"], pred["ref_descriptors1"] = ref_desc[0][:, None], ref_desc[1][:, None]
    else:  # early exit: the reference's torch.stack([]) has nothing to return (A12 ii)
        pred["ref_descriptors0"] = desc0.new_zeros((1, 0, m, D))
        pred["ref_descriptors1"] = desc1.new_zeros((1, 0, n, D))
    return pred
