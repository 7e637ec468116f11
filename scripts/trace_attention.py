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
names = ["wg0 pre", "wg0 gotS", "wg0 arr", "wg1 pre", "wg1 gotS", "wg1 arr"]
nb = int((tr[0] > 0).sum())
for j in range(min(nb, 56)):
    print(f"blk={j:3d} " + "  ".join(f"{names[s]}={tr[s, j] - t0:6d}" for s in range(6)) +
          f" | mma wake(t0,t1)={tr[6, 2 * j] - t0:6d},{tr[6, 2 * j + 1] - t0:6d} issued={tr[7, 2 * j] - t0:6d},{tr[7, 2 * j + 1] - t0:6d}")

print("item epilogues (per item: start, pv_done, o_free, stored) wg0 | wg1")
for i in range(0, nb, 16):
    print(f"kv_base={i:3d} wg0 " + " ".join(f"{tr[8 + k, i] - t0:6d}" for k in range(4)) + " | wg1 " + " ".join(f"{tr[12 + k, i] - t0:6d}" for k in range(4)))

if os.environ.get("LG_TRACE_SPLIT"):
    print("split softmax, tile 0 half 0 warp 0, cycles: wait S | sweep | tile barrier | P store + arrive | to next block || tile 1 lag")
    for j in range(33, 47):
        print(f"blk={j:3d} {tr[1, j] - tr[0, j]:5d} {tr[16, j] - tr[1, j]:5d} {tr[17, j] - tr[16, j]:5d} {tr[2, j] - tr[17, j]:5d} {tr[0, j + 1] - tr[2, j]:5d}"
              f" || {tr[5, j] - tr[2, j]:6d}   period {tr[0, j + 1] - tr[0, j]:5d}")
elif (tr[16] > 0).any():
    print("wg0 warp 0 inside a block, cycles after gotS: first chunk loaded | groups 0-2 | 3-6 | 7-10 | 11-15 + sums | P st issued | "
          "st waited + fenced | bar.sync + arrive | (pre of the next block)")
    for j in range(33, 47):
        t = [tr[1, j]] + [tr[16 + k, j] for k in range(7)] + [tr[2, j], tr[0, j + 1]]
        print(f"blk={j:3d} " + " ".join(f"{t[k + 1] - t[k]:5d}" for k in range(len(t) - 1)))

if (tr[24] > 0).any():
    print("PV warp per global step G (cycles): loop top -> P(G) ready -> PV issued -> commits done | period | P arrived after loop top | "
          "similarity warp: S(G+3) issued after this step's loop top")
    for G in range(70, 96):
        t = [tr[24, G], tr[6, G], tr[26, G], tr[7, G]]
        print(f"G={G:3d} t={G & 1} " + " ".join(f"{t[k + 1] - t[k]:5d}" for k in range(3)) + f" | {tr[24, G + 1] - tr[24, G]:5d}"
              f" | {tr[(G & 1) * 3 + 2, G >> 1] - tr[24, G]:6d} | {tr[27, G + 3] - tr[24, G]:6d}")
