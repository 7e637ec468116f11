"""Assignment + filter at the bench shape (32 pairs x 2048 x 2048): TF32x3 tile kernels vs the split-bf16 strip
kernels (the matcher's path).  ncu: -k regex:assign_strip -c 2 captures pass 1 and pass 2 of the first strip call."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
B, N = 32, 2048
g = torch.Generator().manual_seed(0)
md0, md1 = torch.randn(B, N, 256, generator=g).cuda(), torch.randn(B, N, 256, generator=g).cuda()
z0, z1 = torch.randn(B, N, generator=g).cuda(), torch.randn(B, N, generator=g).cuda()
tp = C.tiles_pad(N)
a, b = C.rows_to_tiles(md0, tp), C.rows_to_tiles(md1, tp)


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


IMPL = "tmem" if os.environ.get("LG_PROFILE_IMPL") == "tmem" else None  # ncu: -k regex:assign_tmem for this variant
print(f"assign_tiles + keys        {timeit(lambda: C.assign_tiles(a, b, tp, tp, z0, z1, 0.1, impl=IMPL)):.3f} ms")
print(f"assign_tiles               {timeit(lambda: C.assign_tiles(a, b, tp, tp, z0, z1, impl=IMPL)):.3f} ms")
if os.environ.get("LG_PROFILE_TF32"):
    s = C.assign(md0, md1, z0, z1)
    print(f"assign (tf32x3)            {timeit(lambda: C.assign(md0, md1, z0, z1)):.3f} ms")
    print(f"filter_matches             {timeit(lambda: C.filter_matches(s, 0.1)):.3f} ms")
    print(f"assign_and_filter (tf32x3) {timeit(lambda: C.assign_and_filter(md0, md1, z0, z1, 0.1)):.3f} ms")
