"""Attention kernel alone on the bench shape (32 pairs x 2048 kpts x 4 heads), for ncu captures and
for CUDA-event timing of kernel variants.  usage: profile_attention.py [B] [N] [reps]"""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
g = torch.Generator().manual_seed(0)
qk = torch.randn(B, N, 256, generator=g).cuda()
v = torch.randn(B, N, 256, generator=g).cuda()
qt, vt = C.pack_cross_qkv(qk, v, None, 4)
tp = C.tiles_pad(N)
out = C.alloc_split(2 * B * tp, 256, "cuda")
qt1, vt1 = C.pack_cross_qkv(v, qk, None, 4)   # the second image of the pair: its own operand tiles
# LG_PROFILE_LENS=n: per-image valid counts (all n) -- the variable-length path of the kernel
L = int(os.environ.get("LG_PROFILE_LENS", 0))
lens = torch.full((B,), L, dtype=torch.int32).cuda() if L else None
s0 = (qt, qt, vt, 0, N, N, lens, lens)
s1 = (qt1, qt1, vt1, B * tp, N, N, lens, lens)


def launch():  # the layer's launch: both images of a SelfBlock in one kernel
    C.attention_tiles2(s0, s1, out, 0.125)


for _ in range(2):
    launch()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    launch()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
fl = 2 * 4.0 * 64 * N * N * B * 4
print(f"attention (both images, one launch) B={B} N={N} lens={L or None}: {ms:.4f} ms/launch, {fl / ms / 1e9:.1f} TFLOP/s")
