"""GPU parity tests, kernel by kernel, through the C ABI (ctypes) against the oracle
(oracle/lightglue_oracle.py, the CPU restatement pinned to the reference module)."""
import numpy as np
import pytest
import torch

from conftest import load_golden, pkg
from oracle import lightglue_oracle as O

pytestmark = pytest.mark.gpu

W = pkg("weights")


@pytest.fixture(scope="module")
def C():
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    return pkg("_cabi")


def _gen(seed):
    g = torch.Generator().manual_seed(seed)
    return g


# ----------------------------------------------------------------------------- (4) filter_matches
def test_filter_known_answers(C):
    """Bit-exact indices on the reference's own score tensors (golden KATs incl. ties, -inf rows)."""
    meta, g = load_golden("filter_matches_kat")
    for name, th in meta["thresholds"].items():
        s = torch.from_numpy(g[name + "_scores"]).cuda().contiguous()
        m0, m1, ms0, ms1 = C.filter_matches(s, th)
        np.testing.assert_array_equal(m0.cpu().numpy(), g[name + "_m0"], err_msg=name)
        np.testing.assert_array_equal(m1.cpu().numpy(), g[name + "_m1"], err_msg=name)
        # expf (GPU libm) vs SLEEF exp (CPU): <= 2 ulp
        np.testing.assert_allclose(ms0.cpu().numpy(), g[name + "_ms0"], rtol=3e-7, atol=0, err_msg=name)
        np.testing.assert_allclose(ms1.cpu().numpy(), g[name + "_ms1"], rtol=3e-7, atol=0, err_msg=name)
        assert m0.dtype == torch.int64 and ms0.dtype == torch.float32


@pytest.mark.parametrize("B,M,N,th", [(1, 1, 1, 0.0), (2, 33, 1000, 0.1), (3, 513, 257, 0.0), (2, 2048, 2048, 0.1),
                                      (1, 4096, 3000, 0.05)])
def test_filter_random_vs_oracle(C, B, M, N, th):
    g = _gen(M * 7 + N)
    # scores quantised to a coarse grid => many exact ties, exercising the lowest-index rule
    s = (torch.randn(B, M + 1, N + 1, generator=g) * 3).round() / 2 - 6
    m0, m1, ms0, ms1 = C.filter_matches(s.cuda(), th)
    r0, r1, rs0, rs1 = O.filter_matches(s, th)
    assert torch.equal(m0.cpu(), r0) and torch.equal(m1.cpu(), r1)
    np.testing.assert_allclose(ms0.cpu().numpy(), rs0.numpy(), rtol=3e-7)
    np.testing.assert_allclose(ms1.cpu().numpy(), rs1.numpy(), rtol=3e-7)


def test_filter_on_reference_log_assignment(C):
    """filter_matches fed the reference's own log_assignment (golden forward) -> its own matches."""
    for name in ("sp_b1_192", "sp_b2_160x224", "adaptive_depth_300x400"):
        meta, g = load_golden("fwd_" + name)
        s = torch.from_numpy(g["log_assignment"]).cuda()
        m0, m1, ms0, ms1 = C.filter_matches(s, meta["conf"]["filter_threshold"])
        np.testing.assert_array_equal(m0.cpu().numpy(), g["matches0"])
        np.testing.assert_array_equal(m1.cpu().numpy(), g["matches1"])
        np.testing.assert_allclose(ms0.cpu().numpy(), g["matching_scores0"], rtol=3e-7)


def test_filter_varlen(C):
    g = _gen(5)
    B, M, N = 3, 300, 260
    s = torch.randn(B, M + 1, N + 1, generator=g)
    ml, nl = [300, 17, 129], [260, 260, 1]
    m0, m1, ms0, ms1 = C.filter_matches(s.cuda(), 0.0, torch.tensor(ml, dtype=torch.int32).cuda(),
                                        torch.tensor(nl, dtype=torch.int32).cuda())
    for b in range(B):
        sub = torch.full((1, ml[b] + 1, nl[b] + 1), 0.0)
        sub[0, :ml[b], :nl[b]] = s[b, :ml[b], :nl[b]]
        r0, r1, rs0, rs1 = O.filter_matches(sub, 0.0)
        assert torch.equal(m0[b, :ml[b]].cpu(), r0[0]) and torch.equal(m1[b, :nl[b]].cpu(), r1[0])
        assert (m0[b, ml[b]:] == -1).all() and (ms1[b, nl[b]:] == 0).all()


# ----------------------------------------------------------------------------- (1) posenc
@pytest.mark.parametrize("B,N,size", [(2, 1000, (640, 480)), (1, 2048, (1600, 1200)), (3, 77, None)])
def test_posenc_vs_oracle(C, B, N, size):
    sd = W.make_state_dict({}, seed=0)
    data = W.make_inputs(B, N, N, seed=3, image_size=size or (640, 480))
    k = data["keypoints0"]
    sz = None if size is None else data["view0"]["image_size"]
    kn = O.normalize_keypoints(k, sz)
    cos_r, sin_r = O.posenc_table(sd, kn, float(N))
    szg = None if sz is None else sz.float().cuda()
    cos, sin = C.posenc(k.cuda(), szg, sd["posenc.Wr.weight"].cuda(), sd["posenc.condition_modulation.weight"].cuda().reshape(-1).contiguous(),
                        sd["posenc.condition_modulation.bias"].cuda(), float(N))
    # phases are O(N) rad: one fp32 ulp of the phase is 2.4e-4 at 2048.  The kernel mirrors the
    # CPU rounding sequence, so the tables agree to sincosf-vs-SLEEF accuracy.
    assert (cos.cpu() - cos_r).abs().max() < 2e-6
    assert (sin.cpu() - sin_r).abs().max() < 2e-6


@pytest.mark.parametrize("with_size", [True, False])
def test_posenc_per_image_condition_vs_oracle(C, with_size):
    """lg_posenc_per_image + lg_keypoint_extent_valid: every image of a padded batch gets the table the oracle computes
    for that image ALONE (its own count as the condition, lightglue.py:490-491; its own extent when size is None,
    :25-26) -- bit for bit what the scalar-condition kernel gives on the un-padded image."""
    sd = W.make_state_dict({}, seed=0)
    wr = sd["posenc.Wr.weight"].cuda()
    wc = sd["posenc.condition_modulation.weight"].cuda().reshape(-1).contiguous()
    bc = sd["posenc.condition_modulation.bias"].cuda()
    lens = [700, 33, 1, 512]
    N = 768
    k = W.make_inputs(4, N, N, seed=8, image_size=(640, 480))["keypoints0"]
    for b, n in enumerate(lens):
        k[b, n:] = 0  # zero padding, as pack_pairs writes it
    size = torch.tensor([[640., 480.]] * 4).cuda() if with_size else None
    ln = torch.tensor(lens, dtype=torch.int32).cuda()
    cos, sin = C.posenc(k.cuda(), size, wr, wc, bc, ln.float(), lens=ln)
    for b, n in enumerate(lens):
        kb = k[b:b + 1, :n]
        kn = O.normalize_keypoints(kb, torch.tensor([[640., 480.]]) if with_size else None)
        cos_r, sin_r = O.posenc_table(sd, kn, float(n))
        # one fp32 ulp of a phase of ~n rad (6e-5 at 700): the CPU's (n,2)x(2,32) GEMV may round mul+add or fma
        assert (cos[b, :n].cpu() - cos_r[0]).abs().max() < 1.3e-4 and (sin[b, :n].cpu() - sin_r[0]).abs().max() < 1.3e-4
        c1, s1 = C.posenc(kb.cuda().contiguous(), None if size is None else size[:1], wr, wc, bc, float(n))
        assert torch.equal(c1[0], cos[b, :n]) and torch.equal(s1[0], sin[b, :n])


# ----------------------------------------------------------------------------- (2) attention
def _unswizzle_tiles(t, n):
    """Inverse of the SWIZZLE_128B tile image: (B,H,T,16384) u8 -> (B,H,T*128,64) bf16 rows."""
    B, H, T, _ = t.shape
    x = t.view(B, H, T, 128, 8, 16)  # row, chunk position, 16 bytes
    r = torch.arange(128, device=t.device) % 8
    pos = (torch.arange(8, device=t.device)[None, :] ^ r[:, None])  # logical chunk c lives at c ^ (r%8)
    x = torch.gather(x, 4, pos[None, None, None, :, :, None].expand(B, H, T, 128, 8, 16))
    return x.reshape(B, H, T * 128, 128).view(torch.bfloat16)[:, :, :n]


def test_pack_self_layout(C):
    B, N, H = 2, 300, 4
    g = _gen(11)
    qkv = torch.randn(B, N, H * 192, generator=g)
    sd = W.make_state_dict({}, seed=0)
    kn = torch.rand(B, N, 2, generator=g) * 2 - 1
    cos, sin = O.posenc_table(sd, kn, float(N))
    q, k, v = O.split_qkv(qkv, H)
    q, k = O.apply_rotary(q, cos, sin), O.apply_rotary(k, cos, sin)
    qt, kt, vt = C.pack_self_qkv(qkv.cuda(), cos.cuda(), sin.cuda(), None, H)
    torch.cuda.synchronize()
    assert torch.equal(_unswizzle_tiles(qt, N).float().cpu(), q.bfloat16().float())
    assert torch.equal(_unswizzle_tiles(kt, N).float().cpu(), k.bfloat16().float())
    # V^T tile: 2 atoms of [64 dh][64 points]
    T = vt.shape[2]
    vv = vt.view(B, H, T, 2, 8192)
    x = vv.view(B, H, T, 2, 64, 8, 16)
    r = torch.arange(64, device=vt.device) % 8
    pos = torch.arange(8, device=vt.device)[None, :] ^ r[:, None]
    x = torch.gather(x, 5, pos[None, None, None, None, :, :, None].expand(B, H, T, 2, 64, 8, 16))
    x = x.reshape(B, H, T, 2, 64, 128).view(torch.bfloat16)           # (.., atom, dh, 64 points)
    x = x.permute(0, 1, 2, 3, 5, 4).reshape(B, H, T * 128, 64)[:, :, :N]
    assert torch.equal(x.float().cpu(), v.bfloat16().float())
    # padded rows are zero
    assert _unswizzle_tiles(qt, T * 128)[:, :, N:].float().abs().max() == 0


def _attn_case(C, B, H, nq, nk, qlen, klen, seed, scale_mag=1.0, check=None):
    g = _gen(seed)
    qk_q = torch.randn(B, nq, H * 64, generator=g) * scale_mag
    v_q = torch.randn(B, nq, H * 64, generator=g)
    qk_k = torch.randn(B, nk, H * 64, generator=g) * scale_mag
    v_k = torch.randn(B, nk, H * 64, generator=g)
    ql = None if qlen is None else torch.tensor(qlen, dtype=torch.int32).cuda()
    kl = None if klen is None else torch.tensor(klen, dtype=torch.int32).cuda()
    qt, _ = C.pack_cross_qkv(qk_q.cuda(), v_q.cuda(), ql, H)
    kt, vt = C.pack_cross_qkv(qk_k.cuda(), v_k.cuda(), kl, H)
    out = C.attention(qt, kt, vt, nq, nk, ql, kl, 0.125)
    torch.cuda.synchronize()
    out = out.cpu()

    def heads(t):
        return t.unflatten(-1, (H, 64)).transpose(1, 2)

    errs = []
    for b in (range(B) if check is None else check):
        a = nq if qlen is None else qlen[b]
        c = nk if klen is None else klen[b]
        if c == 0 or a == 0:
            assert out[b].abs().max() == 0
            continue
        q, k, v = heads(qk_q[b:b + 1, :a]), heads(qk_k[b:b + 1, :c]), heads(v_k[b:b + 1, :c])
        ref = O.attention(q, k, v).transpose(1, 2).flatten(-2)                     # fp32 oracle
        refb = O.attention(q.bfloat16().float(), k.bfloat16().float(), v.bfloat16().float()).transpose(1, 2).flatten(-2)
        got = out[b:b + 1, :a]
        errs.append(((got - ref).abs().max().item(), (got - refb).abs().max().item()))
        assert out[b, a:].abs().max() == 0 if a < nq else True
    return errs


@pytest.mark.parametrize("B,nq,nk", [(1, 128, 128), (1, 256, 256), (2, 300, 200), (1, 1024, 1024), (2, 2048, 2048),
                                     (1, 97, 4096)])
def test_attention_vs_oracle(C, B, nq, nk):
    errs = _attn_case(C, B, 4, nq, nk, None, None, seed=nq + nk)
    for e32, eb in errs:
        # bf16 operands (Q,K,V,P), fp32 accumulate: vs the fp32 oracle ~1e-2 on O(1) outputs; vs the
        # oracle fed bf16-rounded operands only the P rounding remains
        assert e32 < 3e-2 and eb < 1.5e-2, (e32, eb)


def test_attention_large_logits(C):
    """peaked softmax (|logit| ~ 30): exercises the lazy-rescale path."""
    errs = _attn_case(C, 1, 4, 512, 1024, None, None, seed=5, scale_mag=4.0)
    for e32, eb in errs:
        assert eb < 3e-2, (e32, eb)


def test_attention_huge_logits(C):
    """logits jumping by > 2^60 between key blocks: exercises the overflow detection + rescale/redo path."""
    errs = _attn_case(C, 1, 4, 384, 1024, None, None, seed=6, scale_mag=9.0)
    for e32, eb in errs:
        assert eb < 3e-2, (e32, eb)


def test_attention_varlen(C):
    errs = _attn_case(C, 4, 4, 512, 640, [512, 1, 200, 0], [640, 129, 128, 5], seed=9)
    for e32, eb in errs:
        assert e32 < 3e-2 and eb < 1.5e-2, (e32, eb)
    errs = _attn_case(C, 2, 4, 256, 256, [256, 256], [0, 256], seed=10)


@pytest.mark.parametrize("qlen", ["1792", "1100", "random"])
def test_attention_varlen_many_items_per_cta(C, qlen):
    """1024 work items on 148 persistent CTAs with skipped query blocks in between (valid counts below the padded
    length): at item boundaries the MMA warp decodes the next items (per-image lengths = dependent global loads) while
    the softmax warpgroups run ahead over the similarity lookahead.  With ONE P-ready barrier per tile this dead-locked
    (two arrivals before the parity wait; seen as a launch failure through the spin-limit trap)."""
    B, N = 16, 2048
    rng = np.random.default_rng(11)
    ql = [int(qlen)] * B if qlen != "random" else rng.integers(1024, N + 1, B).tolist()
    kl = ql if qlen != "1792" else None  # "1792": key side at full length, the shape of the first reproducer
    errs = _attn_case(C, B, 4, N, N, ql, kl, seed=21, check=[0, 7, 15])
    for e32, eb in errs:
        assert e32 < 3e-2 and eb < 1.5e-2, (e32, eb)


# ----------------------------------------------------------------------------- (3) assignment
@pytest.mark.parametrize("B,M,N", [(1, 128, 128), (2, 200, 333), (1, 1024, 1024), (2, 2048, 2048), (1, 1, 700)])
def test_assign_vs_oracle(C, B, M, N):
    g = _gen(M + 3 * N)
    md0 = torch.randn(B, M, 256, generator=g) * 1.5
    md1 = torch.randn(B, N, 256, generator=g) * 1.5
    z0, z1 = torch.randn(B, M, 1, generator=g) * 2, torch.randn(B, N, 1, generator=g) * 2
    sim = torch.einsum("bmd,bnd->bmn", (md0 / 4).double(), (md1 / 4).double())
    ref64 = O.sigmoid_log_double_softmax(sim, z0.double(), z1.double())
    ref32 = O.sigmoid_log_double_softmax(torch.einsum("bmd,bnd->bmn", md0 / 4, md1 / 4), z0, z1)
    got = C.assign(md0.cuda(), md1.cuda(), z0[..., 0].contiguous().cuda(), z1[..., 0].contiguous().cuda()).cpu()
    e64 = (got.double() - ref64).abs().max().item()
    e32ref = (ref32.double() - ref64).abs().max().item()
    # TF32x3 must be as good as the fp32 reference itself (north_star: 1e-4 abs end to end)
    assert e64 < max(3e-5, 3 * e32ref), (e64, e32ref)
    assert got[:, -1, -1].abs().max() == 0


def _split_tiles(C, md):
    B, N, _ = md.shape
    tp = C.tiles_pad(N)
    return C.rows_to_tiles(md.cuda().contiguous(), tp), tp


def _f16_tiles(C, md):
    """fp32 rows (B, N, 256) -> (fp16 operand tiles of lg_assign_tiles, tiles per image, the fp16-rounded rows).
    Layout built here on the host, independently of the kernels: [row tile][4 K atoms][128 rows][8 chunks of 8 halves],
    16-byte chunk c of row r stored at chunk c ^ (r & 7) (SWIZZLE_128B)."""
    B, N, K = md.shape
    assert K == 256
    tp = C.tiles_pad(N)
    h = torch.zeros(B, tp * 128, K, dtype=torch.float16)
    h[:, :N] = md.half()
    t = h.view(B, tp, 128, 4, 8, 8).permute(0, 1, 3, 2, 4, 5).contiguous()   # (B, tp, atom, row, chunk, 8)
    rows = torch.arange(128).view(128, 1)
    src = torch.arange(8).view(1, 8) ^ (rows & 7)                           # stored chunk p holds logical chunk p ^ (r & 7)
    t = torch.gather(t, 4, src.view(1, 1, 1, 128, 8, 1).expand(B, tp, 4, 128, 8, 8).contiguous())
    tiles = t.view(torch.uint8).reshape(B * tp, 4, 16384).cuda()
    return tiles, tp, h[:, :N].float()


@pytest.mark.parametrize("B,M,N", [(1, 128, 128), (2, 200, 333), (1, 1024, 1024), (2, 2048, 2048), (1, 1, 700), (3, 129, 1),
                                   (2, 300, 2049)])
def test_assign_tiles_vs_oracle(C, B, M, N):
    """lg_assign_tiles (fp16 operand tiles, the matcher's path) against the fp64 evaluation of
    sigmoid_log_double_softmax (lightglue.py:284-296) on the SAME fp16-rounded operands: what is left is the fp32
    accumulation and the exp2 / log2 arithmetic (bound 2e-4 on |sim| ~ 10); the rounding of the operands themselves is
    reported against the unrounded fp64 result (bound 1e-2, inside the bf16 path's 2e-2)."""
    g = _gen(M + 3 * N)
    md0 = torch.randn(B, M, 256, generator=g) * 1.5
    md1 = torch.randn(B, N, 256, generator=g) * 1.5
    z0, z1 = torch.randn(B, M, 1, generator=g) * 2, torch.randn(B, N, 1, generator=g) * 2
    a, tpa, r0 = _f16_tiles(C, md0)
    b, tpb, r1 = _f16_tiles(C, md1)
    ref64 = O.sigmoid_log_double_softmax(torch.einsum("bmd,bnd->bmn", (r0 / 4).double(), (r1 / 4).double()), z0.double(), z1.double())
    exact = O.sigmoid_log_double_softmax(torch.einsum("bmd,bnd->bmn", (md0 / 4).double(), (md1 / 4).double()), z0.double(), z1.double())
    got = C.assign_tiles(a, b, tpa, tpb, z0[..., 0].contiguous().cuda(), z1[..., 0].contiguous().cuda()).cpu()
    assert got.shape == (B, M + 1, N + 1) and torch.isfinite(got).all()
    err, err_exact = (got.double() - ref64).abs().max().item(), (got.double() - exact).abs().max().item()
    print(f"assign_tiles {B}x{M}x{N}: {err:.2e} vs fp64 on the rounded operands, {err_exact:.2e} vs unrounded")
    assert err < 2e-4 and err_exact < 1e-2
    assert got[:, -1, -1].abs().max() == 0


def test_linear_f16_tiles_match_host_layout(C):
    """lg_linear_f16 (final_proj's epilogue) writes exactly the tile image _f16_tiles builds on the host from the fp32
    result of the same linear: the two sides of the operand format agree byte for byte up to fp16 rounding ties."""
    g = _gen(3)
    B, N = 2, 300
    x = torch.randn(B, N, 256, generator=g)
    w = torch.randn(256, 256, generator=g) / 16
    bias = torch.randn(256, generator=g)
    xt, tp = _split_tiles(C, x)
    pw = C.pack_weight(w.cuda().contiguous())
    got = C.linear_f16(xt, pw, bias.cuda(), 256)
    want_rows = C.linear_f32(xt, pw, bias.cuda(), 256).view(B, tp * 128, 256)[:, :N].cpu()
    want, _, _ = _f16_tiles(C, want_rows)
    gh, wh = got.view(torch.float16).float().cpu(), want.view(torch.float16).float().cpu()
    valid = torch.zeros(B, tp * 128, dtype=torch.bool)
    valid[:, :N] = True
    valid = valid.view(B * tp, 1, 128, 1).expand(B * tp, 4, 128, 64).reshape(B * tp, 4, 8192)
    assert (gh - wh)[valid].abs().max() <= 2e-3 * wh.abs().max()   # one fp16 ulp of the largest value
    assert ((gh - wh)[valid] != 0).float().mean() < 1e-3           # and only on rounding ties


@pytest.mark.parametrize("B,M,N,varlen", [(1, 300, 200, False), (2, 257, 640, True), (1, 128, 128, False),
                                          (2, 97, 130, True), (3, 1500, 1100, True), (12, 1610, 1580, True)])
def test_assign_tiles_keys_equal_filter_scan(C, B, M, N, varlen):
    """keys gathered in pass 2 of lg_assign_tiles + lg_filter_from_keys == lg_filter_matches (one scan of the matrix)
    on the same scores, bit for bit, including ties and variable-length batches; padding is -inf."""
    g = torch.Generator().manual_seed(B * 1000 + M + N)
    md0 = torch.randn(B, M, 256, generator=g)
    md1 = torch.randn(B, N, 256, generator=g)
    md1[:, 5] = md1[:, 3]
    md0[:, 7] = md0[:, 2]
    md1[:, : min(M, N) // 2] += 2.0 * md0[:, : min(M, N) // 2]
    z0, z1 = torch.randn(B, M, generator=g), torch.randn(B, N, generator=g)
    z1[:, 5] = z1[:, 3]
    z0[:, 7] = z0[:, 2]
    # (12 pairs: several items per persistent CTA with one or two valid row tiles and skipped items in between -- the
    # accumulator hand-shake of the row-tile-1 warpgroups must survive that)
    mls = ([M, max(1, M - 70), max(1, M // 3)] + [max(1, (M * (3 + 5 * i)) // 53) for i in range(9)])[:B]
    nls = ([max(1, N - 150), N, max(1, N // 5)] + [max(1, N - 131 * i) for i in range(9)])[:B]
    ml = torch.tensor(mls, dtype=torch.int32).cuda() if varlen else None
    nl = torch.tensor(nls, dtype=torch.int32).cuda() if varlen else None
    a, tpa, r0 = _f16_tiles(C, md0)
    b, tpb, r1 = _f16_tiles(C, md1)
    s_plain = C.assign_tiles(a, b, tpa, tpb, z0.cuda(), z1.cuda(), None, ml, nl)
    s_keys, f_keys = C.assign_tiles(a, b, tpa, tpb, z0.cuda(), z1.cuda(), 0.05, ml, nl)
    assert torch.equal(s_plain, s_keys)
    f_scan = C.filter_matches(s_keys, 0.05, ml, nl)
    for x, y in zip(f_scan, f_keys):
        assert torch.equal(x, y)
    assert (f_keys[0] >= 0).sum() > 0
    # against the TF32x3 tile kernel on the same (fp16-rounded) operands: same padding pattern, values within fp32 noise
    s_ref = C.assign(r0.cuda(), r1.cuda(), z0.cuda(), z1.cuda(), ml, nl)
    fin = torch.isfinite(s_ref)
    assert torch.equal(fin, torch.isfinite(s_keys))
    assert (s_ref[fin] - s_keys[fin]).abs().max() < 3e-4


def test_logit_rows(C):
    g = _gen(5)
    x, w, b = torch.randn(3, 77, 256, generator=g), torch.randn(256, generator=g), torch.randn(1, generator=g)
    got = C.logit_rows(x.cuda(), w.cuda(), b.cuda()).cpu()
    assert got.shape == (3, 77)
    assert (got - (x @ w + b)).abs().max() < 2e-5


def test_assign_varlen(C):
    g = _gen(77)
    B, M, N = 3, 300, 260
    ml, nl = [300, 17, 129], [260, 260, 1]
    md0, md1 = torch.randn(B, M, 256, generator=g), torch.randn(B, N, 256, generator=g)
    z0, z1 = torch.randn(B, M, 1, generator=g), torch.randn(B, N, 1, generator=g)
    got = C.assign(md0.cuda(), md1.cuda(), z0[..., 0].contiguous().cuda(), z1[..., 0].contiguous().cuda(),
                   torch.tensor(ml, dtype=torch.int32).cuda(), torch.tensor(nl, dtype=torch.int32).cuda()).cpu()
    for b in range(B):
        a, c = ml[b], nl[b]
        sim = torch.einsum("bmd,bnd->bmn", md0[b:b + 1, :a] / 4, md1[b:b + 1, :c] / 4)
        ref = O.sigmoid_log_double_softmax(sim, z0[b:b + 1, :a], z1[b:b + 1, :c])[0]
        assert (got[b, :a, :c] - ref[:a, :c]).abs().max() < 3e-5
        assert (got[b, :a, N] - ref[:a, c]).abs().max() < 1e-6
        assert (got[b, M, :c] - ref[a, :c]).abs().max() < 1e-6
        if a < M:
            assert torch.isinf(got[b, a:M, :N]).all()


# ----------------------------------------------------------------------------- (5) confidence / pruning
@pytest.mark.parametrize("B,N,use_conf", [(1, 1024, True), (3, 777, True), (2, 4096, False), (1, 5, True)])
def test_prune_plan_bit_exact_on_reference_scores(C, B, N, use_conf):
    """Keep mask + order-preserving compaction + stop count, fed the oracle's own score tensors."""
    g = _gen(N + B)
    conf = torch.rand(B, N, generator=g) if use_conf else None
    match = torch.rand(B, N, generator=g)
    # plant exact-threshold values: `<=` / `<` / `>` boundaries must behave like torch
    thr = O.confidence_threshold(2, 9)
    wc = 0.95
    match[:, ::7] = float(np.float32(1 - wc))
    if use_conf:
        conf[:, ::5] = float(np.float32(thr))
    keep_idx, counts = C.prune_plan(None if conf is None else conf.cuda(), match.cuda(), thr, float(np.float32(1 - wc)))
    for b in range(B):
        mask = O.pruning_mask(None if conf is None else conf[b:b + 1], match[b:b + 1], thr, wc)
        want = torch.where(mask)[1]
        k = int(counts[b, 0])
        assert k == want.numel()
        assert torch.equal(keep_idx[b, :k].cpu().long(), want)
        if use_conf:
            assert int(counts[b, 1]) == int((conf[b] < thr).sum())


@pytest.mark.parametrize("name", ["fwd_adaptive_both_2048", "trace_adaptive_both_512x400", "trace_adaptive_width_300x400"])
def test_prune_plan_bit_exact_on_the_reference_run_tensors(C, name):
    """north_star: pruning indices bit-exact when fed the reference's OWN score tensors -- the token confidences
    (lightglue.py:102-107), matchability scores (:317-318) and keep masks (:586-593) the reference module produced at
    every pruning layer of an adaptive run (hooks in oracle/make_golden_scale.py), not synthetic ones."""
    from conftest import load_golden

    meta, g = load_golden(name)
    width = meta["conf"]["width_confidence"]
    min_match = float(np.float32(1 - width))
    assert meta["layers_pruned"] > 0
    for i in range(meta["layers_pruned"]):
        thr = float(meta["confidence_thresholds"][i])
        for side in (0, 1):
            match = torch.from_numpy(g[f"L{i}_match{side}"]).cuda()
            conf = torch.from_numpy(g[f"L{i}_conf{side}"]).cuda() if f"L{i}_conf{side}" in g else None
            keep = torch.from_numpy(g[f"L{i}_keep{side}"])
            keep_idx, counts = C.prune_plan(conf, match, thr, min_match)
            want = torch.where(keep)[1]
            k = int(counts[0, 0])
            assert k == want.numel(), (name, i, side)
            assert torch.equal(keep_idx[0, :k].cpu().long(), want), (name, i, side)
            if conf is not None:  # the count check_if_stop needs (:595-606)
                assert int(counts[0, 1]) == int((conf < thr).sum())


def test_token_scores_and_gather(C):
    g = _gen(3)
    B, N = 2, 1500
    sd = W.make_state_dict({}, seed=0, adaptive_fixture=True)
    desc = torch.randn(B, N, 256, generator=g)
    conf_r, _ = O.token_confidence(sd, 3, desc, desc)
    match_r = O.matchability(sd, 3, desc)
    conf, match = C.token_scores(desc.cuda(), sd["token_confidence.3.token.0.weight"].cuda().reshape(-1),
                                 sd["token_confidence.3.token.0.bias"].cuda(),
                                 sd["log_assignment.3.matchability.weight"].cuda().reshape(-1),
                                 sd["log_assignment.3.matchability.bias"].cuda())
    assert (conf.cpu() - conf_r).abs().max() < 2e-6 and (match.cpu() - match_r).abs().max() < 2e-6
    # gather
    cos, sin = torch.randn(B, N, 32, generator=g), torch.randn(B, N, 32, generator=g)
    ind = torch.stack([torch.randperm(2 * N, generator=g)[:N].sort().values for _ in range(B)])
    prune = torch.ones(B, 2 * N, dtype=torch.int64)
    keep_idx, counts = C.prune_plan(conf, match, 0.85, 0.05)
    d_o, c_o, s_o, i_o = C.prune_gather(desc.cuda(), cos.cuda(), sin.cuda(), ind.cuda(), keep_idx, counts, pr := prune.cuda())
    for b in range(B):
        k = int(counts[b, 0])
        sel = keep_idx[b, :k].cpu().long()
        assert torch.equal(d_o[b, :k].cpu(), desc[b].index_select(0, sel))
        assert torch.equal(c_o[b, :k].cpu(), cos[b].index_select(0, sel))
        assert torch.equal(s_o[b, :k].cpu(), sin[b].index_select(0, sel))
        assert torch.equal(i_o[b, :k].cpu(), ind[b].index_select(0, sel))
        want = prune[b].clone()
        want[ind[b].index_select(0, sel)] += 1
        assert torch.equal(pr[b].cpu(), want)


@pytest.mark.parametrize("B,M,N,varlen,use_conf", [(2, 1500, 1100, False, True), (3, 700, 2048, True, True), (1, 129, 5, False, False)])
def test_score_prune_tiles_equals_the_row_kernels(C, B, M, N, varlen, use_conf):
    """lg_score_prune_tiles (scores + keep mask + scan + compaction in one launch, directly on split tiles) against the
    chain it replaces -- tiles_to_rows, lg_token_scores, lg_prune_plan, lg_prune_gather, rows_to_tiles -- and with it
    the oracle's mask: kept counts, unconfident counts, original indices, prune counters and the compacted descriptor /
    rotary rows, image by image; rows behind the kept ones are zero."""
    g = _gen(B + M + N)
    sd = W.make_state_dict({}, seed=0, adaptive_fixture=True)
    wt = sd["token_confidence.3.token.0.weight"].cuda().reshape(-1) if use_conf else None
    bt = sd["token_confidence.3.token.0.bias"].cuda() if use_conf else None
    wm, bm = sd["log_assignment.3.matchability.weight"].cuda().reshape(-1), sd["log_assignment.3.matchability.bias"].cuda()
    thr, min_match = 0.85, float(np.float32(0.05))
    tp = [C.tiles_pad(M), C.tiles_pad(N)]
    n_cur, n_orig = [M, N], [2 * M, 2 * N]
    lens = [torch.tensor([M, max(1, M // 3), max(1, M - 77)][:B], dtype=torch.int32).cuda(),
            torch.tensor([max(1, N - 131), N, 1][:B], dtype=torch.int32).cuda()] if varlen else [None, None]
    desc = [torch.randn(B, n, 256, generator=g).cuda() for n in n_cur]
    cos = [torch.randn(B, n, 32, generator=g).cuda() for n in n_cur]
    sin = [torch.randn(B, n, 32, generator=g).cuda() for n in n_cur]
    ind = [torch.stack([torch.randperm(no, generator=g)[:n].sort().values for _ in range(B)]).cuda()
           for n, no in zip(n_cur, n_orig)]
    prune = [torch.randint(1, 4, (B, no), generator=g).cuda() for no in n_orig]
    rt0 = B * tp[0]
    x = C.alloc_split(B * (tp[0] + tp[1]), 256, "cuda")
    rot = torch.empty(B * (tp[0] + tp[1]), C.SPLIT_ATOM, device="cuda", dtype=torch.uint8)
    C.rows_to_tiles(desc[0], tp[0], out=x[:rt0]); C.rows_to_tiles(desc[1], tp[1], out=x[rt0:])
    C.rotary_tiles(cos[0], sin[0], tp[0], out=rot[:rt0]); C.rotary_tiles(cos[1], sin[1], tp[1], out=rot[rt0:])
    x_out, rot_out = torch.full_like(x, 0x7f), torch.full_like(rot, 0x7f)  # NaN-ish bytes: zero fill must cover them
    ind_out = [torch.full_like(t, -5) for t in ind]
    prune_out = [torch.zeros_like(t) for t in prune]
    lens_out = [torch.zeros(B, dtype=torch.int32, device="cuda") for _ in range(2)]
    counts = torch.zeros(B, 4, dtype=torch.int32, device="cuda")
    C.score_prune_tiles(x, rot, x_out, rot_out, ind, ind_out, prune, prune_out, lens, lens_out, counts, wt, bt, wm, bm,
                        thr, min_match, True, B, M, N, tp[0], tp[1], n_orig[0], n_orig[1])
    rows_out = [C.tiles_to_rows(x_out[:rt0], B, tp[0] * 128, 256, tp[0]), C.tiles_to_rows(x_out[rt0:], B, tp[1] * 128, 256, tp[1])]
    for side in range(2):
        n = n_cur[side]
        rows_in = C.tiles_to_rows(x[:rt0] if side == 0 else x[rt0:], B, n, 256, tp[side])  # what the tiles hold (hi + lo)
        conf, match = C.token_scores(rows_in, wt, bt, wm, bm, lens[side])
        keep_idx, cnt = C.prune_plan(conf, match, thr, min_match, lens[side])
        for b in range(B):
            nv = n if lens[side] is None else int(lens[side][b])
            k = int(cnt[b, 0])
            assert int(counts[b, 2 * side]) == k and int(lens_out[side][b]) == k, (side, b)
            if use_conf:
                assert int(counts[b, 2 * side + 1]) == int(cnt[b, 1])
                mask = O.pruning_mask(conf[b:b + 1, :nv].cpu(), match[b:b + 1, :nv].cpu(), thr, 0.95)
                assert int(mask.sum()) == k
            sel = keep_idx[b, :k].long()
            assert torch.equal(ind_out[side][b, :k], ind[side][b].index_select(0, sel))
            want_pr = prune[side][b].clone()
            want_pr[ind[side][b].index_select(0, sel)] += 1
            assert torch.equal(prune_out[side][b], want_pr)
            assert torch.equal(rows_out[side][b, :k], rows_in[b].index_select(0, sel))
            assert (rows_out[side][b, k:] == 0).all()
        # rotary rows: compare through a second compaction by the row kernels
        d_o, c_o, s_o, _ = C.prune_gather(rows_in.contiguous(), cos[side], sin[side], ind[side], keep_idx, cnt,
                                          prune[side].clone(), zero=True)
        want_rot = C.rotary_tiles(c_o, s_o, tp[side]).view(B, tp[side] * 2, 128, 128)  # (image, tile x {cos, sin}, row, 128 B)
        got_rot = (rot_out[:rt0] if side == 0 else rot_out[rt0:]).view(B, tp[side] * 2, 128, 128)
        for b in range(B):
            k = int(cnt[b, 0])
            rows = torch.arange(tp[side] * 128, device="cuda").view(tp[side], 1, 128).expand(tp[side], 2, 128).reshape(-1, 128)
            live = (rows < k)[:, :, None].expand(-1, -1, 128)
            assert torch.equal(got_rot[b][live], want_rot[b][live])
            assert (got_rot[b][~live] == 0).all()  # padding rows: zeros (finite; the rotary of a padding row is never used)
    assert torch.equal(prune[0], prune[0]) and (x_out.view(torch.int16) != 0x7f7f).any()


@pytest.mark.parametrize("B,M,N,varlen", [(1, 300, 200, False), (2, 257, 640, True), (1, 128, 128, False), (2, 97, 130, True)])
def test_assign_keys_equals_assign_plus_filter(B, M, N, varlen):
    """lg_assign_keys + lg_filter_from_keys == lg_assign + lg_filter_matches, bit for bit (scores and all four
    filter outputs), including ties (duplicated descriptors) and variable-length batches."""
    C = pkg("_cabi")
    g = torch.Generator().manual_seed(B * 1000 + M + N)
    md0 = torch.randn(B, M, 256, generator=g)
    md1 = torch.randn(B, N, 256, generator=g)
    md1[:, 5] = md1[:, 3]            # equal columns -> ties in every row
    md0[:, 7] = md0[:, 2]            # equal rows -> ties in every column
    md1[:, : min(M, N) // 2] += 2.0 * md0[:, : min(M, N) // 2]   # real matches
    z0, z1 = torch.randn(B, M, generator=g), torch.randn(B, N, generator=g)
    z1[:, 5] = z1[:, 3]
    z0[:, 7] = z0[:, 2]
    ml = torch.tensor([M, max(1, M - 70)][:B], dtype=torch.int32).cuda() if varlen else None
    nl = torch.tensor([max(1, N - 150), N][:B], dtype=torch.int32).cuda() if varlen else None
    args = (md0.cuda(), md1.cuda(), z0.cuda(), z1.cuda())
    s_ref = C.assign(*args, ml, nl)
    f_ref = C.filter_matches(s_ref, 0.05, ml, nl)
    s_new, f_new = C.assign_and_filter(*args, 0.05, ml, nl)
    assert torch.equal(s_ref, s_new)
    for a, b in zip(f_ref, f_new):
        assert torch.equal(a, b)
    assert (f_new[0] >= 0).sum() > 0


# ----------------------------------------------------------------------------- out-of-bounds canaries
def _guarded(nbytes, dtype=torch.uint8, fill=0x5A):
    """buffer of nbytes with 64 KB guard zones on both sides (compute-sanitizer is not available on the pool)."""
    g = 65536
    raw = torch.full((nbytes + 2 * g,), fill, dtype=torch.uint8, device="cuda")
    return raw, raw[g:g + nbytes], g


def _guards_intact(raw, g, fill=0x5A):
    return bool((raw[:g] == fill).all()) and bool((raw[-g:] == fill).all())


@pytest.mark.parametrize("B,M,N,varlen", [(2, 200, 333, False), (3, 129, 257, True), (1, 1, 700, False), (2, 1024, 640, True)])
def test_assign_tiles_writes_stay_in_bounds(C, B, M, N, varlen):
    """lg_assign_tiles through the raw C ABI with guard zones around the scores, the workspace and the key buffer; the
    padding columns of the pitched rows (N+1 .. pitch-1) are never written."""
    lib = C.lib()
    g = _gen(B + M + N)
    a, tpa, _ = _f16_tiles(C, torch.randn(B, M, 256, generator=g))
    b, tpb, _ = _f16_tiles(C, torch.randn(B, N, 256, generator=g))
    z0, z1 = torch.randn(B, M, generator=g).cuda(), torch.randn(B, N, generator=g).cuda()
    ml = torch.tensor([M, max(1, M // 2), 1][:B], dtype=torch.int32).cuda() if varlen else None
    nl = torch.tensor([max(1, N - 130), N, 5][:B], dtype=torch.int32).cuda() if varlen else None
    pitch = C.scores_pitch(N)
    raw_s, s, gs = _guarded(B * (M + 1) * pitch * 4)
    s.view(torch.float32).fill_(7.0)
    raw_w, w, gw = _guarded(lib.lg_assign_tiles_ws_bytes(B, M, N))
    raw_k, k, gk = _guarded(lib.lg_filter_ws_bytes(B, M, N))
    rc = lib.lg_assign_tiles(a.data_ptr(), b.data_ptr(), tpa, tpb, z0.data_ptr(), z1.data_ptr(),
                             None if ml is None else ml.data_ptr(), None if nl is None else nl.data_ptr(), 1.0 / 16,
                             s.data_ptr(), pitch, w.data_ptr(), k.data_ptr(), B, M, N, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert rc == 0
    assert _guards_intact(raw_s, gs) and _guards_intact(raw_w, gw) and _guards_intact(raw_k, gk)
    got = s.view(torch.float32).view(B, M + 1, pitch)
    want = C.assign_tiles(a, b, tpa, tpb, z0, z1, None, ml, nl)
    assert torch.equal(got[:, :, : N + 1], want)
    assert (got[:, :, N + 1:] == 7.0).all()


def test_gemm_f32_strided_batches_and_bounds(C):
    """lg_gemm_f32 on strided views with padding columns: values against torch, padding and guard zones untouched."""
    g = _gen(9)
    Bt, M, N, K, ldc = 3, 70, 45, 37, 52
    A = torch.randn(Bt, M, K, generator=g).cuda()
    Bm = torch.randn(Bt, N, K, generator=g).cuda()
    bias = torch.randn(N, generator=g).cuda()
    raw, c, gz = _guarded(Bt * M * ldc * 4)
    out = c.view(torch.float32).view(Bt, M, ldc)
    out.fill_(7.0)
    C.gemm_f32(A, Bm, out, M, N, K, K, K, ldc, bias=bias, alpha=0.5, batch=Bt, inner=1,
               strides=(M * K, 0, N * K, 0, M * ldc, 0, 0, 0))
    torch.cuda.synchronize()
    want = 0.5 * torch.einsum("bmk,bnk->bmn", A.double(), Bm.double()) + bias.double()
    assert (out[:, :, :N].double() - want).abs().max() < 1e-4
    assert (out[:, :, N:] == 7.0).all() and _guards_intact(raw, gz)
    # (K, N) form of B with a valid-K count per batch
    Bk = torch.randn(Bt, K, N, generator=g).cuda()
    klen = torch.tensor([K, 5, 20], dtype=torch.int32).cuda()
    out.fill_(7.0)
    C.gemm_f32(A, Bk, out, M, N, K, K, N, ldc, b_kn=True, batch=Bt, inner=1,
               strides=(M * K, 0, K * N, 0, M * ldc, 0, 0, 0), k_len=klen)
    for b in range(Bt):
        kk = int(klen[b])
        want = A[b, :, :kk].double() @ Bk[b, :kk].double()
        assert (out[b, :, :N].double() - want).abs().max() < 1e-4
    assert _guards_intact(raw, gz)
