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

    def forward(self, desc0, desc1, cs0, cs1, len0, len1, n_layers):
        B, m, n = self.B, self.m, self.n
        side0, side1 = (0, m), (self.T0, n)
        if self.w.input_proj is None:
            self.x[: self.T0].view(B, m, D).copy_(desc0)
            self.x[self.T0:].view(B, n, D).copy_(desc1)
        else:
            wi, bi = self.w.input_proj
            raw = torch.cat([desc0.reshape(self.T0, -1), desc1.reshape(self.T - self.T0, -1)], 0).contiguous()
            d_in = raw.shape[1]
            C.gemm_f32(raw, wi, self.x, self.T, D, d_in, d_in, d_in, D, bias=bi)
        cos = torch.cat([cs0[0].reshape(-1, 32), cs1[0].reshape(-1, 32)], 0).contiguous()
        sin = torch.cat([cs0[1].reshape(-1, 32), cs1[1].reshape(-1, 32)], 0).contiguous()
        for i in range(n_layers):
            blk_s, blk_c = self.w.layers[i]
            # SelfBlock (lightglue.py:178-191): q | k | v blocks of 256 columns, head h at h*64
            self._linear(self.x, blk_s["wqkv"], blk_s["bqkv"], self.qkv, n_out=3 * D, k=D, lda=D, ldb=D, ldc=3 * D)
            C.rotary_rows_f32(self.qkv.view(1, self.T, 3 * D), cos, sin, 0, D, H)
            self._attend(side0, side0, 0, D, 2 * D, (len0, len0))
            self._attend(side1, side1, 0, D, 2 * D, (len1, len1))
            self._ffn(blk_s)
            # CrossBlock (lightglue.py:220-249): shared to_qk projection (columns 0..255), to_v (256..511)
            self._linear(self.x, blk_c["wqkv"], blk_c["bqkv"], self.qkv, n_out=2 * D, k=D, lda=D, ldb=D, ldc=3 * D)
            self._attend(side0, side1, 0, 0, D, (len0, len1))
            self._attend(side1, side0, 0, 0, D, (len1, len0))
            self._ffn(blk_c)
        return self.x[: self.T0].view(B, m, D), self.x[self.T0:].view(B, n, D)

    def assign(self, i, th, len0, len1):
        """MatchAssignment (lightglue.py:306-315) + filter_matches (:321-337): fp32 final_proj, TF32x3 similarity."""
        wf, bf, wm, bm = self.w.final[i]
        md = torch.empty(self.T, D, device=self.dev, dtype=torch.float32)
        self._linear(self.x, wf, bf, md, n_out=D, k=D, lda=D, ldb=D, ldc=D)
        z = C.logit_rows(self.x.view(1, self.T, D), wm, bm)[0]
        B, m, n = self.B, self.m, self.n
        return C.assign_and_filter(md[: self.T0].view(B, m, D), md[self.T0:].view(B, n, D), z[: self.T0].view(B, m),
                                   z[self.T0:].view(B, n), th, len0, len1, prescale=D ** -0.25)
