"""HBM-bound passes on 128 images x 2048 tokens (4x the bench batch, so that the inputs exceed the L2): achieved GB/s against the algorithmic bytes of
SURVEY.md 8(d) -- confidence / matchability scores, prune plan, order-preserving gather, matchability logits,
tile <-> row format changes, filter from keys."""
import importlib, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
C = importlib.import_module("cs566-project-lightglue_b200._cabi")
B, N = 128, 2048  # 268 MB of descriptors: larger than the 126 MB L2, every timed launch streams from HBM
g = torch.Generator().manual_seed(0)
desc = torch.randn(B, N, 256, generator=g).cuda()
cos, sin = torch.randn(B, N, 32, generator=g).cuda(), torch.randn(B, N, 32, generator=g).cuda()
w1, w2 = torch.randn(256, generator=g).cuda(), torch.randn(256, generator=g).cuda()
b1, b2 = torch.zeros(1).cuda(), torch.zeros(1).cuda()
ind = torch.arange(N).cuda()[None].repeat(B, 1)
prune = torch.ones(B, N, dtype=torch.int64).cuda()
peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] \
    if os.path.exists("MEASURED_PEAKS.json") else 6550.0


def timeit(fn, reps=20):
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


def report(name, ms, nbytes):
    gbs = nbytes / ms / 1e6
    print(f"{name:34s} {ms * 1e3:8.1f} us  {nbytes / 1e6:8.1f} MB  {gbs:7.0f} GB/s  {gbs / peak:5.1%} of {peak:.0f}")


conf, match = C.token_scores(desc, w1, b1, w2, b2)
keep, cnt = C.prune_plan(conf, match, 0.85, 0.4)
kept = int(cnt[:, 0].sum())
print(f"kept {kept} of {B * N} tokens")
report("token_scores (conf + matchability)", timeit(lambda: C.token_scores(desc, w1, b1, w2, b2)), B * N * (1024 + 8))
report("prune_plan", timeit(lambda: C.prune_plan(conf, match, 0.85, 0.4)), B * N * 12)
report("prune_gather", timeit(lambda: C.prune_gather(desc, cos, sin, ind, keep, cnt, prune)), kept * 2 * (1024 + 256 + 8) + kept * 20)
report("logit_rows", timeit(lambda: C.logit_rows(desc, w1, b1)), B * N * (1024 + 4))
tp = C.tiles_pad(N)
t = C.rows_to_tiles(desc, tp)
report("rows_to_tiles", timeit(lambda: C.rows_to_tiles(desc, tp, out=t)), B * N * 2048)
report("tiles_to_rows", timeit(lambda: C.tiles_to_rows(t, B, N, 256, tp)), B * N * 2048)
