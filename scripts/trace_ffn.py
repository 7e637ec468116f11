"""Debug: per-tile timeline (SM cycles) of one CTA of the fused FFN kernel (build with make EXTRA=-DLG_FFN_TRACE)."""
import ctypes, importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
B, N = 32, 2048
rt = 2 * B * C.tiles_pad(N)
g = torch.Generator().manual_seed(0)
x = C.alloc_split(rt, 256, "cuda"); x.random_(0, 255)
ctx = C.alloc_split(rt, 256, "cuda"); ctx.random_(0, 255)
h = C.alloc_split(rt, 512, "cuda")
w0 = C.pack_weight(torch.randn(512, 512, generator=g).cuda() * 0.05)
b = torch.zeros(512, device="cuda"); gamma = torch.ones(512, device="cuda")
for _ in range(3):
    C.ffn_ln_gelu(x, w0, b, gamma, b, 1e-5, a1=ctx, out=h)
torch.cuda.synchronize()
buf = np.zeros(16 * 64, dtype=np.uint64)
lib = C.lib()
lib.lg_debug_ffn_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert lib.lg_debug_ffn_trace(buf.ctypes.data, 16 * 64) == 0
tr = buf.reshape(16, 64).astype(np.int64)
t0 = tr[tr > 0].min()
names = ["mma start", "mma end", "epi wait acc", "acc ready", "stats sent", "peer stats", "atom0", "atom1", "atom2", "atom3"]
n = int((tr[0] > 0).sum())
for i in range(n):
    print(f"tile {i:2d} " + "  ".join(f"{names[s]}={tr[s, i] - t0:7d}" for s in range(10)))
