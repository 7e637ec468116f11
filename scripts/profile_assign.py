"""Assignment + filter at the bench shape (32 pairs x 2048 x 2048): the fp16-tile kernels of the matcher's path
(lg_assign_f16.cu) and, with LG_PROFILE_TF32=1, the TF32x3 kernels of the fp32 path.
ncu: -k regex:assign_f16 -c 4 captures stats<rows>, stats<cols>, merge is a separate kernel, write."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
g = torch.Generator().manual_seed(0)
x0, x1 = torch.randn(B, N, 256, generator=g).cuda(), torch.randn(B, N, 256, generator=g).cuda()
z0, z1 = torch.randn(B, N, generator=g).cuda(), torch.randn(B, N, generator=g).cuda()
tp = C.tiles_pad(N)
w = C.pack_weight((torch.randn(256, 256, generator=g) / 16).cuda())
bias = torch.zeros(256).cuda()
a = C.linear_f16(C.rows_to_tiles(x0, tp), w, bias, 256)  # fp16 operand tiles, as final_proj writes them
b = C.linear_f16(C.rows_to_tiles(x1, tp), w, bias, 256)


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


t_keys = timeit(lambda: C.assign_tiles(a, b, tp, tp, z0, z1, 0.1))
t_plain = timeit(lambda: C.assign_tiles(a, b, tp, tp, z0, z1))
gb = B * (N + 1) * (N + 1) * 4 / 1e9
print(f"assign_tiles + keys        {t_keys:.3f} ms   ({gb / t_keys * 1e3:.0f} GB/s of log_assignment written)")
print(f"assign_tiles               {t_plain:.3f} ms   ({gb / t_plain * 1e3:.0f} GB/s)")
if os.environ.get("LG_PROFILE_TF32"):
    md0, md1 = x0, x1
    s = C.assign(md0, md1, z0, z1)
    print(f"assign (tf32x3)            {timeit(lambda: C.assign(md0, md1, z0, z1)):.3f} ms")
    print(f"filter_matches             {timeit(lambda: C.filter_matches(s, 0.1)):.3f} ms")
    print(f"assign_and_filter (tf32x3) {timeit(lambda: C.assign_and_filter(md0, md1, z0, z1, 0.1)):.3f} ms")
