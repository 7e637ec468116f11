"""Debug: per-iteration timeline of one attention CTA (build with make EXTRA=-DLG_ATTN_TRACE)."""
import ctypes, importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
B, N = 32, 2048
g = torch.Generator().manual_seed(0)
qk = torch.randn(B, N, 256, generator=g).cuda(); v = torch.randn(B, N, 256, generator=g).cuda()
qt, vt = C.pack_cross_qkv(qk, v, None, 4)
out = C.alloc_split(B * C.tiles_pad(N), 256, "cuda")
for _ in range(3):
    C.attention_tiles(qt, qt, vt, out, 0, N, N, None, None, 0.125)
torch.cuda.synchronize()
buf = np.zeros(8192, dtype=np.uint64)
lib = C.lib()
lib.lg_debug_attn_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert lib.lg_debug_attn_trace(buf.ctypes.data, 8192) == 0
tr = buf.reshape(32, 256).astype(np.int64)
t0 = tr[tr > 0].min()
names = ["mma wake", "mma issued", "wg0 pre-wait", "wg0 got S", "wg0 arrived", "wg1 pre-wait", "wg1 got S", "wg1 arrived"]
for j in range(16):
    print(f"j={j:2d} " + "  ".join(f"{names[s][:11]}={tr[s, j] - t0:6d}" for s in range(2, 8)))
print("mma (n, wake, issued):")
for n in range(32):
    print(f"  n={n:2d} wake={tr[0, n] - t0:6d} issued={tr[1, n] - t0:6d}")
