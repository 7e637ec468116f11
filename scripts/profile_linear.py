"""The four per-token linear launches of one block on the bench shape (32 pairs x 2 x 2048 rows), timed alone with
CUDA events.  LG_LINEAR_PAIR=0 selects the single-CTA kernels.  usage: profile_linear.py [B] [N] [reps]"""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dev = "cuda"
tp = C.tiles_pad(N)
rt = 2 * B * tp
g = torch.Generator().manual_seed(0)
x = C.alloc_split(rt, 256, dev); x.random_(0, 255)
ctx = C.alloc_split(rt, 256, dev); ctx.random_(0, 255)
h = C.alloc_split(rt, 512, dev); h.random_(0, 255)
x2 = C.alloc_split(rt, 256, dev)
hpre = torch.empty(rt * 128, 512, device=dev)
w0 = C.pack_weight(torch.randn(512, 512, generator=g).cuda() * 0.05)
w3 = C.pack_weight(torch.randn(256, 512, generator=g).cuda() * 0.05)
wq = C.pack_weight(torch.randn(768, 256, generator=g).cuda() * 0.05, 192)
wc = C.pack_weight(torch.randn(512, 256, generator=g).cuda() * 0.05)
b512, b256, b768 = torch.zeros(512, device=dev), torch.zeros(256, device=dev), torch.zeros(768, device=dev)
rot = torch.zeros(rt, C.SPLIT_ATOM, device=dev, dtype=torch.uint8)
qkv = [tuple(C.alloc_tiles(B, 4, N, dev) for _ in range(3)) for _ in range(2)]
(q0, k0, v0), (q1, k1, v1) = qkv
gamma, beta = torch.ones(512, device=dev), torch.zeros(512, device=dev)
cases = {
    "ffn0  linear_f32   K512 N512": (lambda: C.linear_f32(x, w0, b512, 512, a1=ctx, out=hpre), 2.0 * rt * 128 * 512 * 512 * 3),
    "ffn3  linear_split K512 N256": (lambda: C.linear_split(h, w3, b256, 256, resid=x, out=x2), 2.0 * rt * 128 * 512 * 256 * 3),
    "qkv_self           K256 N768": (lambda: C.linear_qkv_self(x, wq, b768, rot, B, tp, tp, qkv), 2.0 * rt * 128 * 256 * 768 * 3),
    "qkv_cross          K256 N512": (lambda: C.linear_qkv_cross(x, wc, b512, B, tp, tp, ((q0, v0), (q1, v1))), 2.0 * rt * 128 * 256 * 512 * 3),
    "final linear_f32   K256 N256": (lambda: C.linear_f32(x, w3[: 256 * 256 * 4], b256, 256), 2.0 * rt * 128 * 256 * 256 * 3),
    "final linear_f16   K256 N256": (lambda: C.linear_f16(x, w3[: 256 * 256 * 4], b256, 256), 2.0 * rt * 128 * 256 * 256 * 3),
    "ffn0+LN+GELU fused K512 N512": (lambda: C.ffn_ln_gelu(x, w0, b512, gamma, beta, 1e-5, a1=ctx, out=h), 2.0 * rt * 128 * 512 * 512 * 3),
    "ln_gelu": (lambda: C.ln_gelu_tiles(hpre, gamma, beta, 1e-5, out=h), 0.0),
}
for name, (fn, fl) in cases.items():
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{name}: {ms * 1e3:8.1f} us/launch  {fl / ms / 1e9:8.1f} TFLOP/s (3-term issued)")
