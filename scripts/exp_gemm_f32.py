"""lg_gemm_f32 per shape of the fp32 path (LG_GEMM_F32_TILE=small|big forces the 64x64 / 128xBN kernel)."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


print("tile =", os.environ.get("LG_GEMM_F32_TILE", "auto"))
for B in (1, 4):
    T = B * 4096
    for n, k in ((768, 256), (512, 256), (256, 256), (256, 512)):
        a, w, c = torch.randn(T, k).cuda(), torch.randn(n, k).cuda(), torch.empty(T, n).cuda()
        us = timeit(lambda: C.gemm_f32(a, w, c, T, n, k, k, k, n))
        print(f"B={B} linear {T}x{n}x{k}: {us:7.1f} us  {2 * T * n * k / us / 1e6:5.1f} TF/s")
    nq = nk = 2048
    H = 4
    q, kk = torch.randn(B * H, nq, 64).cuda(), torch.randn(B * H, nk, 64).cuda()
    s = torch.empty(B * H, nq, nk).cuda()
    us = timeit(lambda: C.gemm_f32(q, kk, s, nq, nk, 64, 64, 64, nk, batch=B * H, inner=1,
                                   strides=(nq * 64, 0, nk * 64, 0, nq * nk, 0, 0, 0)))
    print(f"B={B} q.k^T  {B * H}x{nq}x{nk}x64: {us:7.1f} us  {2 * B * H * nq * nk * 64 / us / 1e6:5.1f} TF/s")
    v, o = torch.randn(B * H, nk, 64).cuda(), torch.empty(B * H, nq, 64).cuda()
    us = timeit(lambda: C.gemm_f32(s, v, o, nq, 64, nk, nk, 64, 64, b_kn=True, batch=B * H, inner=1,
                                   strides=(nq * nk, 0, nk * 64, 0, nq * 64, 0, 0, 0)))
    print(f"B={B} P.v    {B * H}x{nq}x64x{nk}: {us:7.1f} us  {2 * B * H * nq * nk * 64 / us / 1e6:5.1f} TF/s")
