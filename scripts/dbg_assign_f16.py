import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
import test_kernels_gpu as T
rng = np.random.default_rng(5)
for (B, M, N) in [(4, 1610, 1580), (12, 1610, 1580), (12, 2048, 2048), (12, 1330, 1610)]:
    g = torch.Generator().manual_seed(M + 3 * N)
    md0 = torch.randn(B, M, 256, generator=g)
    md1 = torch.randn(B, N, 256, generator=g)
    z0, z1 = torch.randn(B, M, generator=g).cuda(), torch.randn(B, N, generator=g).cuda()
    a, tpa, r0 = T._f16_tiles(C, md0)
    b, tpb, r1 = T._f16_tiles(C, md1)
    ml = torch.tensor(rng.integers(400, M + 1, B), dtype=torch.int32).cuda()
    nl = torch.tensor(rng.integers(400, N + 1, B), dtype=torch.int32).cuda()
    print(B, M, N, ml.tolist(), nl.tolist(), flush=True)
    s, f = C.assign_tiles(a, b, tpa, tpb, z0, z1, 0.05, ml, nl)
    torch.cuda.synchronize()
    f2 = C.filter_matches(s, 0.05, ml, nl)
    print("   ok, keys == scan:", all(torch.equal(x, y) for x, y in zip(f, f2)), flush=True)
