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


REPS = int(os.environ.get("LG_PROFILE_REPS", 20))  # 1 under ncu: one warm-up + one timed launch per kernel


def timeit(fn, reps=REPS):
    for _ in range(3 if reps > 1 else 1):
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

# the adaptive step of the matcher's path: ONE fused launch on the split tiles (lg_score_prune_tiles): 64 pairs x 2 images
Bp = B // 2
x = torch.cat([C.rows_to_tiles(desc[:Bp], tp), C.rows_to_tiles(desc[Bp:], tp)], 0)
rot = torch.cat([C.rotary_tiles(cos[:Bp], sin[:Bp], tp), C.rotary_tiles(cos[Bp:], sin[Bp:], tp)], 0)
x2, rot2 = torch.empty_like(x), torch.empty_like(rot)
i0 = ind[:Bp].contiguous(); i1 = ind[Bp:].contiguous()
i0n, i1n = torch.empty_like(i0), torch.empty_like(i1)
p0 = prune[:Bp].contiguous(); p1 = prune[Bp:].contiguous()
p0n, p1n = torch.empty_like(p0), torch.empty_like(p1)
l0 = torch.full((Bp,), N, dtype=torch.int32).cuda(); l1 = l0.clone()
l0n, l1n = torch.empty_like(l0), torch.empty_like(l1)
counts = torch.empty(Bp, 4, dtype=torch.int32).cuda()
fused = lambda: C.score_prune_tiles(x, rot, x2, rot2, (i0, i1), (i0n, i1n), (p0, p1), (p0n, p1n), (l0, l1), (l0n, l1n), counts,
                                    w1, b1, w2, b2, 0.85, 0.4, True, Bp, N, N, tp, tp, N, N)
fused()
kept2 = int(counts[:, 0].sum() + counts[:, 2].sum())
# reads every row (1 KB) ; kept rows written (1 KB + 256 B rotary + 8 B index) and their rotary rows read; counters 16 B / token
report("score_prune_tiles (fused, kept %d)" % kept2, timeit(fused), B * N * 1024 + kept2 * (1024 + 512 + 16) + B * N * 16)
scores_only = lambda: C.score_prune_tiles(x, None, None, None, None, None, None, None, (l0, l1), None, counts, w1, b1, w2, b2,
                                          0.85, 0.4, False, Bp, N, N, tp, tp, N, N)
report("score_prune_tiles (depth only)", timeit(scores_only), B * N * 1024)
