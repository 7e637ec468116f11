"""B200-native NearestNeighborMatcher forward -- drop-in for the reference's
gluefactory/models/matchers/nearest_neighbor_matcher.py:42-72: same conf keys (`ratio_thresh`,
`distance_thresh`, `mutual_check`, `loss`), same data dict in (`descriptors0/1`), same pred dict out (`matches0/1`,
`matching_scores0/1`, `similarity`, `log_assignment`).  The similarity GEMM + double log-softmax run on the
assignment kernels of the LightGlue path (TF32x3 tcgen05, csrc/lg_assign.cu), find_nn + mutual_check in
csrc/lg_nn.cu.  Select it with `model.matcher.name: b200glue.nn_matcher`.  Inference only (the N-pair loss is
training code, out of scope); CUDA only, no fallback.
"""
import torch
from torch import nn

from . import _cabi as C
from .matcher import merge_conf


class NearestNeighborMatcher(nn.Module):
    default_conf = {
        "name": "nearest_neighbor_matcher",
        "ratio_thresh": None,
        "distance_thresh": None,
        "mutual_check": True,
        "loss": None,
    }
    required_data_keys = ["descriptors0", "descriptors1"]

    def __init__(self, conf=None) -> None:
        super().__init__()
        self.conf = merge_conf(self.default_conf, conf)
        if self.conf.loss is not None:
            raise NotImplementedError("NearestNeighborMatcher: training losses are out of scope (inference only)")
        self.eval()

    @torch.no_grad()
    def forward(self, data: dict) -> dict:
        for key in self.required_data_keys:
            assert key in data, f"Missing key {key} in data"
        d0, d1 = data["descriptors0"], data["descriptors1"]
        if not d0.is_cuda:
            raise RuntimeError("b200glue NearestNeighborMatcher runs on CUDA (sm_100a) only; move the batch to the GPU")
        d0, d1 = d0.contiguous().float(), d1.contiguous().float()
        D = d0.shape[-1]
        if D != d1.shape[-1] or D > 256:
            raise ValueError(f"NearestNeighborMatcher: descriptor width {D} / {d1.shape[-1]}: the similarity kernel takes "
                             "equal widths up to 256")
        if D < 256:  # the reference takes any width (its shipped NN configs use 128-d DISK / ALIKED descriptors):
            # zero columns do not change a dot product
            d0 = torch.nn.functional.pad(d0, (0, 256 - D))
            d1 = torch.nn.functional.pad(d1, (0, 256 - D))
        with torch.cuda.device(d0.device):
            return self._match(data, d0, d1)

    def _match(self, data, d0, d1):
        lens = [None if data.get(k) is None else torch.as_tensor(data[k], device=d0.device).to(torch.int32).contiguous()
                for k in ("num_keypoints0", "num_keypoints1")]
        sim, la = C.nn_assign(d0, d1, *lens)
        m0, m1, ms0, ms1 = C.nn_match(sim, self.conf.ratio_thresh, self.conf.distance_thresh, self.conf.mutual_check, *lens)
        return {"matches0": m0, "matches1": m1, "matching_scores0": ms0, "matching_scores1": ms1,
                "similarity": sim, "log_assignment": la}


__main_model__ = NearestNeighborMatcher
