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
out = C.alloc_split(B * tp, 256, "cuda")
for _ in range(2):
    C.attention_tiles(qt, qt, vt, out, 0, N, N, None, None, 0.125)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    C.attention_tiles(qt, qt, vt, out, 0, N, N, None, None, 0.125)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
fl = 4.0 * 64 * N * N * B * 4
print(f"attention B={B} N={N}: {ms:.4f} ms/launch, {fl / ms / 1e9:.1f} TFLOP/s")
