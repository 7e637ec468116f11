"""Assignment + filter at the bench shape (32 pairs x 2048 x 2048): separate launches vs the fused keys path."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
B, N = 32, 2048
g = torch.Generator().manual_seed(0)
md0, md1 = torch.randn(B, N, 256, generator=g).cuda(), torch.randn(B, N, 256, generator=g).cuda()
z0, z1 = torch.randn(B, N, generator=g).cuda(), torch.randn(B, N, generator=g).cuda()


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


s = C.assign(md0, md1, z0, z1)
print(f"assign                {timeit(lambda: C.assign(md0, md1, z0, z1)):.3f} ms")
print(f"filter_matches        {timeit(lambda: C.filter_matches(s, 0.1)):.3f} ms")
print(f"assign_and_filter     {timeit(lambda: C.assign_and_filter(md0, md1, z0, z1, 0.1)):.3f} ms")
