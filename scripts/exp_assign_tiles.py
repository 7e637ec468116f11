"""lg_assign_tiles (split-bf16 strips, A resident) vs lg_assign (TF32x3 tiles): accuracy against fp64 and time at the
bench shape (32 pairs x 2048 x 2048)."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
from oracle import lightglue_oracle as O


def tiles(md):
    B, N, _ = md.shape
    tp = C.tiles_pad(N)
    t = C.alloc_split(B * tp, 256, md.device)
    C.rows_to_tiles(md, tp, out=t)
    return t, tp


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for B, M, N in [(1, 128, 128), (2, 200, 333), (1, 1024, 1024), (2, 2048, 2048), (1, 1, 700)]:
    g = torch.Generator().manual_seed(M + 3 * N)
    md0 = torch.randn(B, M, 256, generator=g) * 1.5
    md1 = torch.randn(B, N, 256, generator=g) * 1.5
    z0, z1 = torch.randn(B, M, generator=g) * 2, torch.randn(B, N, generator=g) * 2
    sim = torch.einsum("bmd,bnd->bmn", (md0 / 4).double(), (md1 / 4).double())
    ref64 = O.sigmoid_log_double_softmax(sim, z0[..., None].double(), z1[..., None].double())
    a, tpa = tiles(md0.cuda())
    b, tpb = tiles(md1.cuda())
    got = C.assign_tiles(a, b, tpa, tpb, z0.cuda(), z1.cuda()).cpu()
    old = C.assign(md0.cuda(), md1.cuda(), z0.cuda(), z1.cuda()).cpu()
    print(f"B={B} M={M} N={N}: |strip - f64| = {(got.double() - ref64).abs().max():.3e}   |tf32x3 - f64| = "
          f"{(old.double() - ref64).abs().max():.3e}", flush=True)

B, N = 32, 2048
g = torch.Generator().manual_seed(0)
md0, md1 = torch.randn(B, N, 256, generator=g).cuda(), torch.randn(B, N, 256, generator=g).cuda()
z0, z1 = torch.randn(B, N, generator=g).cuda(), torch.randn(B, N, generator=g).cuda()
a, tpa = tiles(md0)
b, tpb = tiles(md1)
print(f"assign (tf32x3)            {timeit(lambda: C.assign(md0, md1, z0, z1)):.3f} ms")
print(f"assign_and_filter (tf32x3) {timeit(lambda: C.assign_and_filter(md0, md1, z0, z1, 0.1)):.3f} ms")
print(f"assign_tiles               {timeit(lambda: C.assign_tiles(a, b, tpa, tpb, z0, z1)):.3f} ms")
print(f"assign_tiles + keys        {timeit(lambda: C.assign_tiles(a, b, tpa, tpb, z0, z1, 0.1)):.3f} ms")
s, f = C.assign_tiles(a, b, tpa, tpb, z0, z1, 0.1)
f2 = C.filter_matches(s, 0.1)
print("keys == scan:", all(torch.equal(x, y) for x, y in zip(f, f2)))
