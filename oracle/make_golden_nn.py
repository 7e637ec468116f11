"""Golden vectors of the reference NearestNeighborMatcher (run in the build container, /root/reference mounted):
python -m oracle.make_golden_nn  ->  tests/golden/nn_*.npz.  Inputs are regenerated from seeds by the tests."""
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

W = importlib.import_module("cs566-project-lightglue_b200.weights")

CASES = {
    "nn_mutual_b2_300x257": dict(conf={"mutual_check": True}, batch=2, m=300, n=257, seed=21, correlated=0.7),
    "nn_ratio_b1_512x400": dict(conf={"ratio_thresh": 0.9, "mutual_check": True}, batch=1, m=512, n=400, seed=22, correlated=0.3),
    "nn_dist_nomutual_b1_130x200": dict(conf={"distance_thresh": 1.1, "mutual_check": False}, batch=1, m=130, n=200, seed=23,
                                        correlated=0.3),
    "nn_all_b1_1024": dict(conf={"ratio_thresh": 0.92, "distance_thresh": 1.12, "mutual_check": True}, batch=1, m=1024, n=1024,
                           seed=24, correlated=0.35),
}


def main():
    mod = ref_shim.load_reference_nn_module()
    for name, c in CASES.items():
        data = W.make_inputs(c["batch"], c["m"], c["n"], seed=c["seed"], correlated=c["correlated"])
        model = mod.NearestNeighborMatcher(dict(c["conf"])).eval()
        with torch.no_grad():
            pred = model._forward(data)
        meta = {**c, "source": "gluefactory/models/matchers/nearest_neighbor_matcher.py:52-72 (reference run)"}
        out = {"meta": json.dumps(meta)}
        for k in ("matches0", "matches1", "matching_scores0", "matching_scores1"):
            out[k] = pred[k].numpy()
        out["log_assignment_sample"] = pred["log_assignment"][:, ::7, ::5].numpy()
        out["similarity_sample"] = pred["similarity"][:, ::7, ::5].numpy()
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), **out)
        print(name, "matched", int((pred["matches0"] > -1).sum()), "of", pred["matches0"].numel())


if __name__ == "__main__":
    main()
