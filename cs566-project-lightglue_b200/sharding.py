"""Pair-level data parallelism over the GPUs of one box.

Image pairs are independent units of LightGlue.forward (no cross-pair term; the only
batch-coupled quantity, the positional-encoding condition, is the padded keypoint count and is
identical on every rank), so a batch shards by pair index with NO collective in the hot path.
Only the small per-pair results (matches + scores: 2*N*12 bytes per pair) are gathered at the
end; `log_assignment` (16.8 MB/pair at 2048) stays on the producing GPU.
One process per GPU; torch.distributed (NCCL over NVLink on the box, gloo in the CPU tests) is
plumbing only.
"""
import torch
import torch.distributed as dist

RESULT_KEYS = ("matches0", "matches1", "matching_scores0", "matching_scores1")


def shard_range(total_pairs: int, world: int, rank: int):
    """Contiguous block of ceil(total/world) pairs for `rank` (last ranks may get fewer / none)."""
    per = -(-total_pairs // world)
    lo = min(rank * per, total_pairs)
    return lo, min(lo + per, total_pairs)


def shard_batch(data: dict, world: int, rank: int) -> dict:
    """Slice every batched tensor of a reference-style data dict to this rank's pairs."""
    total = data["keypoints0"].shape[0]
    lo, hi = shard_range(total, world, rank)

    def cut(v):
        if isinstance(v, dict):
            return {k: cut(x) for k, x in v.items()}
        if isinstance(v, torch.Tensor) and v.dim() > 0 and v.shape[0] == total:
            return v[lo:hi]
        return v

    return {k: cut(v) for k, v in data.items()}


def gather_results(pred: dict, total_pairs: int, group=None) -> dict:
    """All-gather the small per-pair outputs so every rank holds the full (total_pairs, N) results.
    Ranks hold equal-size blocks except possibly the tail, so blocks are padded to the block size
    for the collective and trimmed afterwards."""
    world = dist.get_world_size(group)
    per = -(-total_pairs // world)
    out = {}
    for k in RESULT_KEYS:
        t = pred[k]
        pad = torch.zeros((per,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
        full = torch.empty((world * per,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(full, pad, group=group)
        out[k] = full[:total_pairs]
    return out
