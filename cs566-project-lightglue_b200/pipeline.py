"""Host-buffer front end of the matcher: the loop of the reference's exporters / evaluators,

    for data in loader:  data = batch_to_device(data, device, non_blocking=True);  pred = model(data);  ... .cpu()

(gluefactory/scripts/export_megadepth.py and gluefactory/eval/eval_pipeline.py drive `TwoViewPipeline` this way,
the matcher being the `forward` of gluefactory/models/matchers/lightglue.py:444-579), restated as a three-stream
pipeline so that the PCIe copies of neighbouring batches hide under the forward of the current one:

    copy-in stream :  H2D(i+1)                       (pinned host -> static device slot)
    compute stream :            forward(i)           (all lg_* kernels, the caller's current stream)
    copy-out stream:                       D2H(i-1)  (matches / scores -> pinned host slot)

Ordering is by CUDA events only; the host blocks once per batch, on the copy-out event of the batch it yields.
"""
import torch

RESULT_KEYS = ("matches0", "matches1", "matching_scores0", "matching_scores1")


def _map(data, fn):
    return {k: (_map(v, fn) if isinstance(v, dict) else fn(v)) for k, v in data.items()}


def _copy_into(dst, src):
    for k, v in src.items():
        if isinstance(v, dict):
            _copy_into(dst[k], v)
        else:
            dst[k].copy_(v, non_blocking=True)


PAIR_KEYS = ("keypoints", "descriptors", "scales", "oris")


def pack_pairs(pairs, length0=None, length1=None, multiple=128):
    """One padded batch out of single-pair data dicts with unequal keypoint counts -- the batching the reference's
    exporter cannot do (utils/export_predictions.py:32-79 feeds one pair per call; its pad helper is
    models/utils/misc.py:19-57 `pad_to_length(mode="zeros")`, mirrored here: zeros behind the valid points).

    pairs: list of reference-style dicts with (1, n_i, .) tensors `keypoints0/1`, `descriptors0/1` (+ `scales*`,
    `oris*` when used) and optional `view0/1.image_size`.  Returns a dict with (B, N_pad, .) tensors, `num_keypoints0/1`
    (B,) int32 and `view*.image_size` (B, 2) f32; N_pad = the longest count rounded up to `multiple` (the kernels' row
    tile), or `length0/1`.  For a pair without `image_size` the reference's size=None rule (1 + max - min over the
    pair's own points, lightglue.py:25-26) is applied here, so the padding never enters it.
    With conf `b200.condition: valid` the matcher then reproduces, pair by pair, the one-pair-per-call results."""
    assert len(pairs) > 0
    out = {}
    for side, length in (("0", length0), ("1", length1)):
        counts = [int(p["keypoints" + side].shape[-2]) for p in pairs]
        n_pad = length if length is not None else max(multiple, -(-max(counts) // multiple) * multiple)
        assert n_pad >= max(counts)
        for key in PAIR_KEYS:
            name = key + side
            if name not in pairs[0]:
                continue
            first = pairs[0][name]
            width = first.shape[-1] if key in ("keypoints", "descriptors") else 1
            stacked = first.new_zeros((len(pairs), n_pad, width))  # one allocation, one slice copy per pair
            for b, (p, c) in enumerate(zip(pairs, counts)):
                stacked[b, :c] = p[name].reshape(c, width)
            out[name] = stacked[..., 0] if key in ("scales", "oris") else stacked
        out["num_keypoints" + side] = torch.tensor(counts, dtype=torch.int32, device=pairs[0]["keypoints" + side].device)
        sizes = []
        for p in pairs:
            v = p.get("view" + side)
            sz = None if v is None else v.get("image_size")
            k = p["keypoints" + side].reshape(-1, 2).float()
            if sz is None:
                sz = 1 + k.max(0).values - k.min(0).values if k.shape[0] > 0 else k.new_ones(2)
            sizes.append(torch.as_tensor(sz).reshape(-1)[-2:].to(device=k.device, dtype=torch.float32))
        out["view" + side] = {"image_size": torch.stack(sizes, 0)}
    return out


def group_by_length(pairs, group: int):
    """Index groups of at most `group` pairs, pairs of similar size together (sorted by the larger keypoint count), so
    that `pack_pairs` pads each group only to ITS longest pair: attention cost grows with the square of the padded
    length.  Returns a list of index lists; results can be put back in order with the indices."""
    order = sorted(range(len(pairs)), key=lambda i: max(int(pairs[i]["keypoints0"].shape[-2]),
                                                        int(pairs[i]["keypoints1"].shape[-2])))
    return [order[g:g + group] for g in range(0, len(order), group)]


def unpack_pred(pred, batch):
    """Per-pair views of a batched prediction, cut to each pair's valid counts (the shapes the one-pair-per-call
    forward returns): list of dicts with `matches0/1`, `matching_scores0/1` (1, n_i) and, when present,
    `log_assignment` (1, m_i + 1, n_i + 1) = the valid block plus the dustbin row / column."""
    l0, l1 = batch["num_keypoints0"].tolist(), batch["num_keypoints1"].tolist()
    outs = []
    for b, (m, n) in enumerate(zip(l0, l1)):
        d = {"matches0": pred["matches0"][b:b + 1, :m], "matches1": pred["matches1"][b:b + 1, :n],
             "matching_scores0": pred["matching_scores0"][b:b + 1, :m],
             "matching_scores1": pred["matching_scores1"][b:b + 1, :n]}
        if "log_assignment" in pred:
            s = pred["log_assignment"][b]
            rows = torch.cat([torch.arange(m, device=s.device), torch.tensor([s.shape[0] - 1], device=s.device)])
            cols = torch.cat([torch.arange(n, device=s.device), torch.tensor([s.shape[1] - 1], device=s.device)])
            d["log_assignment"] = s.index_select(0, rows).index_select(1, cols)[None]
        outs.append(d)
    return outs


class HostPipeline:
    """`for pred in HostPipeline(model).run(batches)`: `batches` yields reference-style data dicts of (pinned) host
    tensors; `pred` holds pinned host tensors for `keys`, valid until `depth` more batches have been yielded (the
    results rotate through depth + 1 pinned slots).  The shape may change between batches (groups of
    `pack_pairs` / `group_by_length` pad to different lengths): the batch in flight is drained and handed out first,
    then the slots are re-allocated for the new shape."""

    def __init__(self, model, depth: int = 2, keys=RESULT_KEYS):
        self.model, self.depth, self.keys = model, max(2, int(depth)), tuple(keys)
        self.dev = next(model.parameters()).device
        if self.dev.type != "cuda":
            raise RuntimeError("HostPipeline needs the model on a CUDA device: there is no CPU path")
        with torch.cuda.device(self.dev):
            self.copy_in = torch.cuda.Stream(self.dev)
            self.copy_out = torch.cuda.Stream(self.dev)
            self.in_ready = [torch.cuda.Event() for _ in range(self.depth)]
            self.fwd_done = [torch.cuda.Event() for _ in range(self.depth)]
            self.out_done = [torch.cuda.Event() for _ in range(self.depth + 1)]
        self.slots = None     # per input slot: device input dict
        self.outs = None      # per output slot (depth + 1 of them): pinned host result dict

    def _alloc(self, batch):
        self.slots = [_map(batch, lambda v: torch.empty(v.shape, dtype=v.dtype, device=self.dev))
                      for _ in range(self.depth)]
        self.shapes = _map(batch, lambda v: (tuple(v.shape), v.dtype))

    def run(self, batches):
        with torch.cuda.device(self.dev):
            yield from self._run(batches)

    def _run(self, batches):
        main = torch.cuda.current_stream(self.dev)
        pending = None  # output slot whose results have been queued but not yet yielded
        i = 0           # batches since the last (re-)allocation: input slot i % depth, output slot i % (depth + 1)
        for batch in batches:
            if self.slots is None or _map(batch, lambda v: (tuple(v.shape), v.dtype)) != self.shapes:
                if pending is not None:  # hand out the batch in flight BEFORE its pinned slot is replaced
                    self.out_done[pending].synchronize()
                    yield self.outs[pending]
                    pending = None
                torch.cuda.synchronize(self.dev)
                self._alloc(batch)
                self.outs = None
                i = 0
            s, o = i % self.depth, i % (self.depth + 1)
            with torch.cuda.stream(self.copy_in):
                if i >= self.depth:
                    self.copy_in.wait_event(self.fwd_done[s])   # forward(i - depth) has consumed this slot
                else:
                    self.copy_in.wait_stream(main)
                _copy_into(self.slots[s], batch)
                self.in_ready[s].record(self.copy_in)
            main.wait_event(self.in_ready[s])
            pred = self.model(self.slots[s])
            self.fwd_done[s].record(main)
            if self.outs is None:
                self.outs = [{k: torch.empty(pred[k].shape, dtype=pred[k].dtype).pin_memory() for k in self.keys}
                             for _ in range(self.depth + 1)]
            with torch.cuda.stream(self.copy_out):
                self.copy_out.wait_event(self.fwd_done[s])
                for k in self.keys:
                    pred[k].record_stream(self.copy_out)
                    self.outs[o][k].copy_(pred[k], non_blocking=True)
                self.out_done[o].record(self.copy_out)
            if pending is not None:
                self.out_done[pending].synchronize()
                yield self.outs[pending]
            pending = o
            i += 1
        if pending is not None:
            self.out_done[pending].synchronize()
            main.wait_stream(self.copy_out)
            yield self.outs[pending]
