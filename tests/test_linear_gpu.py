"""GPU parity tests for the split-tile activations and the tcgen05 per-token linears (lg_linear.cu),
through the C ABI, against torch fp64 / the oracle's functions on the same seeded inputs."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import pkg
from oracle import lightglue_oracle as O

pytestmark = pytest.mark.gpu
W = pkg("weights")


@pytest.fixture(scope="module")
def C():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    return pkg("_cabi")


def _gen(seed):
    return torch.Generator().manual_seed(seed)


def _unswizzle_split(tiles, K):
    """(RT, K/64, 32768) u8 -> hi, lo as (RT*128, K) float32, by plain index arithmetic (no kernel)."""
    RT = tiles.shape[0]
    x = tiles.view(RT, K // 64, 2, 128, 8, 16)
    r = torch.arange(128, device=tiles.device) % 8
    pos = torch.arange(8, device=tiles.device)[None, :] ^ r[:, None]
    x = torch.gather(x, 4, pos[None, None, None, :, :, None].expand(RT, K // 64, 2, 128, 8, 16))
    x = x.reshape(RT, K // 64, 2, 128, 128).view(torch.bfloat16).float()  # (RT, atoms, 2, 128 rows, 64)
    x = x.permute(2, 0, 3, 1, 4).reshape(2, RT * 128, K)
    return x[0], x[1]


def _unswizzle_attn(t, n):
    B, H, T, _ = t.shape
    x = t.view(B, H, T, 128, 8, 16)
    r = torch.arange(128, device=t.device) % 8
    pos = torch.arange(8, device=t.device)[None, :] ^ r[:, None]
    x = torch.gather(x, 4, pos[None, None, None, :, :, None].expand(B, H, T, 128, 8, 16))
    return x.reshape(B, H, T * 128, 128).view(torch.bfloat16)[:, :, :n].float()


def _unswizzle_vt(vt, n):
    B, H, T, _ = vt.shape
    x = vt.view(B, H, T, 2, 64, 8, 16)
    r = torch.arange(64, device=vt.device) % 8
    pos = torch.arange(8, device=vt.device)[None, :] ^ r[:, None]
    x = torch.gather(x, 5, pos[None, None, None, None, :, :, None].expand(B, H, T, 2, 64, 8, 16))
    x = x.reshape(B, H, T, 2, 64, 128).view(torch.bfloat16)
    return x.permute(0, 1, 2, 3, 5, 4).reshape(B, H, T * 128, 64)[:, :, :n].float()


@pytest.mark.parametrize("B,N,K", [(1, 128, 256), (3, 300, 128), (2, 77, 512)])
def test_rows_tiles_roundtrip(C, B, N, K):
    x = torch.randn(B, N, K, generator=_gen(N + K)) * 3
    tp = C.tiles_pad(N)
    t = C.rows_to_tiles(x.cuda(), tp)
    hi, lo = _unswizzle_split(t, K)
    hi, lo = hi.view(B, tp * 128, K), lo.view(B, tp * 128, K)
    want_hi = x.bfloat16().float()
    assert torch.equal(hi[:, :N].cpu(), want_hi)
    assert torch.equal(lo[:, :N].cpu(), (x - want_hi).bfloat16().float())
    assert hi[:, N:].abs().max() == 0 and lo[:, N:].abs().max() == 0
    back = C.tiles_to_rows(t, B, N, K, tp).cpu()
    assert ((back - x).abs() <= x.abs() * 2.0 ** -16).all()


@pytest.mark.parametrize("RT,K0,K1,NOUT", [(1, 256, 0, 256), (5, 256, 256, 512), (3, 128, 0, 256), (150, 512, 0, 256),
                                           (301, 256, 0, 768)])
def test_linear_f32_vs_fp64(C, RT, K0, K1, NOUT):
    g = _gen(RT + K0 + NOUT)
    K = K0 + K1
    x = torch.randn(1, RT * 128, K, generator=g)
    w = (torch.rand(NOUT, K, generator=g) * 2 - 1) / np.sqrt(K)
    b = torch.randn(NOUT, generator=g)
    ref = (x[0].double() @ w.double().t() + b.double())
    a0 = C.rows_to_tiles(x[..., :K0].contiguous().cuda(), RT)
    a1 = C.rows_to_tiles(x[..., K0:].contiguous().cuda(), RT) if K1 else None
    wp = C.pack_weight(w.cuda())
    got = C.linear_f32(a0, wp, b.cuda(), NOUT, a1=a1).cpu().double()
    err = (got - ref).abs().max().item()
    ref32 = (F.linear(x[0], w, b).double() - ref).abs().max().item()
    print(f"RT={RT} K={K} NOUT={NOUT}: split-bf16 err {err:.2e} (fp32 torch err {ref32:.2e})")
    assert err < 4e-5, err  # 3-term split bf16: ~2^-16 relative on O(1) outputs


def test_linear_split_residual_concat(C):
    g = _gen(7)
    RT = 9
    x = torch.randn(1, RT * 128, 512, generator=g)
    msg = torch.randn(1, RT * 128, 512, generator=g)
    res = torch.randn(1, RT * 128, 256, generator=g) * 2
    w = (torch.rand(256, 1024, generator=g) * 2 - 1) / 32
    b = torch.randn(256, generator=g)
    ref = res[0].double() + torch.cat([x, msg], -1)[0].double() @ w.double().t() + b.double()
    a0, a1 = C.rows_to_tiles(x.cuda(), RT), C.rows_to_tiles(msg.cuda(), RT)
    rs = C.rows_to_tiles(res.cuda(), RT)
    out = C.linear_split(a0, C.pack_weight(w.cuda()), b.cuda(), 256, a1=a1, resid=rs)
    got = C.tiles_to_rows(out, 1, RT * 128, 256, RT)[0].cpu().double()
    assert (got - ref).abs().max() < 1e-4
    out2 = C.linear_split(a0, C.pack_weight(w[:, :512].contiguous().cuda()), b.cuda(), 256)
    got2 = C.tiles_to_rows(out2, 1, RT * 128, 256, RT)[0].cpu().double()
    assert (got2 - (x[0].double() @ w[:, :512].double().t() + b.double())).abs().max() < 1e-4


def test_ln_gelu_tiles(C):
    g = _gen(3)
    rows = 5 * 128
    h = torch.randn(rows, 512, generator=g) * 2 + 0.3
    gamma, beta = torch.randn(512, generator=g), torch.randn(512, generator=g)
    ref = F.gelu(F.layer_norm(h, (512,), gamma, beta, 1e-5))
    t = C.ln_gelu_tiles(h.cuda(), gamma.cuda(), beta.cuda(), 1e-5)
    got = C.tiles_to_rows(t, 1, rows, 512, 5)[0].cpu()
    assert (got - ref).abs().max() < 5e-5


@pytest.mark.parametrize("row_tiles", [1, 2, 7, 160])
def test_ffn_ln_gelu_fused(C, row_tiles):
    """lg_ffn_ln_gelu == GELU(LayerNorm([x | ctx] W^T + b)) in fp64, and == the two-kernel path it replaces, on
    rows with very different scales / offsets (exercises the cross-CTA merge of the row statistics)."""
    g = _gen(40 + row_tiles)
    rows = row_tiles * 128
    scale = torch.logspace(-2, 1.5, rows)[torch.randperm(rows, generator=g)][:, None]
    x = torch.randn(rows, 256, generator=g) * scale
    ctx = torch.randn(rows, 256, generator=g) * scale + 0.5 * scale
    w = torch.randn(512, 512, generator=g) * 0.06
    w[:256] += 0.02                      # the two column halves get different means
    b = torch.randn(512, generator=g)
    gamma, beta = torch.randn(512, generator=g), torch.randn(512, generator=g)
    xt = C.rows_to_tiles(x[None].cuda(), row_tiles)
    ct = C.rows_to_tiles(ctx[None].cuda(), row_tiles)
    wp = C.pack_weight(w.cuda())
    got_t = C.ffn_ln_gelu(xt, wp, b.cuda(), gamma.cuda(), beta.cuda(), 1e-5, a1=ct)
    got = C.tiles_to_rows(got_t, 1, rows, 512, row_tiles)[0].cpu()
    # inputs as the kernel sees them (hi + lo of the split tiles), so the reference isolates the kernel's own error
    xs = C.tiles_to_rows(xt, 1, rows, 256, row_tiles)[0].cpu().double()
    cs = C.tiles_to_rows(ct, 1, rows, 256, row_tiles)[0].cpu().double()
    ws = C.tiles_to_rows(C.rows_to_tiles(w[None].cuda(), 4), 1, 512, 512, 4)[0].cpu().double()
    h = torch.cat([xs, cs], 1) @ ws.t() + b.double()
    ref = F.gelu(F.layer_norm(h, (512,), gamma.double(), beta.double(), 1e-5))
    assert (got.double() - ref).abs().max() < 2e-4, (got.double() - ref).abs().max()
    two = C.ln_gelu_tiles(C.linear_f32(xt, wp, b.cuda(), 512, a1=ct), gamma.cuda(), beta.cuda(), 1e-5)
    two = C.tiles_to_rows(two, 1, rows, 512, row_tiles)[0].cpu()
    assert (got - two).abs().max() < 2e-4


def _perm_qkv_self(H=4):
    """reference row h*192 + d*3 + part (lightglue.py:185) -> packed row h*192 + part*64 + d."""
    idx = torch.empty(H * 192, dtype=torch.long)
    for h in range(H):
        for part in range(3):
            for d in range(64):
                idx[h * 192 + part * 64 + d] = h * 192 + d * 3 + part
    return idx


@pytest.mark.parametrize("B,M,N", [(1, 256, 256), (2, 300, 130)])
def test_linear_qkv_self_tiles(C, B, M, N):
    g = _gen(M + N)
    sd = W.make_state_dict({}, seed=0)
    w, bias = sd["transformers.0.self_attn.Wqkv.weight"], sd["transformers.0.self_attn.Wqkv.bias"]
    perm = _perm_qkv_self()
    wp = C.pack_weight(w[perm].contiguous().cuda(), 192)
    bp = bias[perm].contiguous().cuda()
    tp = [C.tiles_pad(M), C.tiles_pad(N)]
    xs = [torch.randn(B, M, 256, generator=g), torch.randn(B, N, 256, generator=g)]
    tabs = [O.posenc_table(sd, torch.rand(B, n, 2, generator=g) * 2 - 1, float(n)) for n in (M, N)]
    a = torch.cat([C.rows_to_tiles(x.cuda(), t) for x, t in zip(xs, tp)], 0)
    rot = torch.cat([C.rotary_tiles(c.cuda(), s.cuda(), t) for (c, s), t in zip(tabs, tp)], 0)
    bufs = [tuple(C.alloc_tiles(B, 4, n, "cuda") for _ in range(3)) for n in (M, N)]
    C.linear_qkv_self(a, wp, bp, rot, B, tp[0], tp[1], bufs)
    torch.cuda.synchronize()
    for side, (x, n) in enumerate(zip(xs, (M, N))):
        q, k, v = O.split_qkv(F.linear(x, w, bias), 4)
        q, k = O.apply_rotary(q, *tabs[side]), O.apply_rotary(k, *tabs[side])
        gq, gk = _unswizzle_attn(bufs[side][0], n).cpu(), _unswizzle_attn(bufs[side][1], n).cpu()
        gv = _unswizzle_vt(bufs[side][2], n).cpu()
        for got, ref in ((gq, q), (gk, k), (gv, v)):
            # bf16 rounding of a value that differs by ~1e-6 from the fp32 reference: <= 1 bf16 ulp
            assert ((got - ref).abs() <= ref.abs() * 2.0 ** -7 + 2e-5).all()
            assert (got == ref.bfloat16().float()).float().mean() > 0.98


def test_linear_qkv_cross_tiles(C):
    g = _gen(5)
    B, M, N = 2, 200, 384
    sd = W.make_state_dict({}, seed=0)
    p = "transformers.2.cross_attn."
    w = torch.cat([sd[p + "to_qk.weight"], sd[p + "to_v.weight"]], 0)
    bias = torch.cat([sd[p + "to_qk.bias"], sd[p + "to_v.bias"]], 0)
    tp = [C.tiles_pad(M), C.tiles_pad(N)]
    xs = [torch.randn(B, M, 256, generator=g), torch.randn(B, N, 256, generator=g)]
    a = torch.cat([C.rows_to_tiles(x.cuda(), t) for x, t in zip(xs, tp)], 0)
    bufs = [tuple(C.alloc_tiles(B, 4, n, "cuda") for _ in range(2)) for n in (M, N)]
    C.linear_qkv_cross(a, C.pack_weight(w.cuda()), bias.cuda(), B, tp[0], tp[1], bufs)
    torch.cuda.synchronize()
    for side, (x, n) in enumerate(zip(xs, (M, N))):
        y = F.linear(x, w, bias)
        qk = y[..., :256].unflatten(-1, (4, 64)).transpose(1, 2)
        v = y[..., 256:].unflatten(-1, (4, 64)).transpose(1, 2)
        for got, ref in ((_unswizzle_attn(bufs[side][0], n).cpu(), qk), (_unswizzle_vt(bufs[side][1], n).cpu(), v)):
            assert ((got - ref).abs() <= ref.abs() * 2.0 ** -7 + 2e-5).all()
            assert (got == ref.bfloat16().float()).float().mean() > 0.98


def test_attention_tiles_matches_rows(C):
    g = _gen(21)
    B, H, nq, nk = 2, 4, 300, 520
    qk_q, v_q = torch.randn(B, nq, 256, generator=g), torch.randn(B, nq, 256, generator=g)
    qk_k, v_k = torch.randn(B, nk, 256, generator=g), torch.randn(B, nk, 256, generator=g)
    ql = torch.tensor([300, 40], dtype=torch.int32).cuda()
    kl = torch.tensor([520, 129], dtype=torch.int32).cuda()
    qt, _ = C.pack_cross_qkv(qk_q.cuda(), v_q.cuda(), ql, H)
    kt, vt = C.pack_cross_qkv(qk_k.cuda(), v_k.cuda(), kl, H)
    ref = C.attention(qt, kt, vt, nq, nk, ql, kl, 0.125)
    tq = C.tiles_pad(nq)
    out = C.alloc_split(3 + B * tq, 256, "cuda")
    C.attention_tiles(qt, kt, vt, out, 3, nq, nk, ql, kl, 0.125)
    got = C.tiles_to_rows(out[3:], B, nq, 256, tq)
    assert ((got - ref).abs() <= ref.abs() * 2.0 ** -16 + 1e-30).all()


def test_linear_hi_only_second_source(C):
    """[a0 | a1] with a1 as plain bf16 tiles (2-term product on those K atoms)."""
    g = _gen(13)
    RT = 6
    x = torch.randn(1, RT * 128, 256, generator=g)
    ctx = torch.randn(1, RT * 128, 256, generator=g).bfloat16().float()
    w = (torch.rand(512, 512, generator=g) * 2 - 1) / 22
    b = torch.randn(512, generator=g)
    ref = torch.cat([x, ctx], -1)[0].double() @ w.double().t() + b.double()
    a0 = C.rows_to_tiles(x.cuda(), RT)
    a1 = C.rows_to_tiles(ctx.cuda(), RT)[:, :, :16384].contiguous()     # hi halves only (lo is exactly zero)
    got = C.linear_f32(a0, C.pack_weight(w.cuda()), b.cuda(), 512, a1=a1, a1_hi_only=True).cpu().double()
    assert (got - ref).abs().max() < 4e-5
