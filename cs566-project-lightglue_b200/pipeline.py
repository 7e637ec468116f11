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


class HostPipeline:
    """`for pred in HostPipeline(model).run(batches)`: `batches` yields reference-style data dicts of (pinned) host
    tensors, all of one shape; `pred` holds pinned host tensors for `keys`, valid until `depth` more batches have
    been yielded."""

    def __init__(self, model, depth: int = 2, keys=RESULT_KEYS):
        self.model, self.depth, self.keys = model, max(2, int(depth)), tuple(keys)
        self.dev = next(model.parameters()).device
        if self.dev.type != "cuda":
            raise RuntimeError("HostPipeline needs the model on a CUDA device: there is no CPU path")
        self.copy_in = torch.cuda.Stream(self.dev)
        self.copy_out = torch.cuda.Stream(self.dev)
        self.slots = None     # per slot: device input dict
        self.outs = None      # per slot: pinned host result dict
        self.in_ready = [torch.cuda.Event() for _ in range(self.depth)]
        self.fwd_done = [torch.cuda.Event() for _ in range(self.depth)]
        self.out_done = [torch.cuda.Event() for _ in range(self.depth)]

    def _alloc(self, batch):
        self.slots = [_map(batch, lambda v: torch.empty(v.shape, dtype=v.dtype, device=self.dev))
                      for _ in range(self.depth)]
        self.shapes = _map(batch, lambda v: (tuple(v.shape), v.dtype))

    def run(self, batches):
        main = torch.cuda.current_stream(self.dev)
        pending = None  # slot whose results have been queued but not yet yielded
        for i, batch in enumerate(batches):
            s = i % self.depth
            if self.slots is None or _map(batch, lambda v: (tuple(v.shape), v.dtype)) != self.shapes:
                torch.cuda.synchronize(self.dev)
                self._alloc(batch)
                self.outs = None
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
                             for _ in range(self.depth)]
            with torch.cuda.stream(self.copy_out):
                self.copy_out.wait_event(self.fwd_done[s])
                for k in self.keys:
                    pred[k].record_stream(self.copy_out)
                    self.outs[s][k].copy_(pred[k], non_blocking=True)
                self.out_done[s].record(self.copy_out)
            if pending is not None:
                self.out_done[pending].synchronize()
                yield self.outs[pending]
            pending = s
        if pending is not None:
            self.out_done[pending].synchronize()
            main.wait_stream(self.copy_out)
            yield self.outs[pending]
