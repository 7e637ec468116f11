"""TEST INFRASTRUCTURE.  Generates tests/golden/*.npz by running the UNMODIFIED reference
module (imported from /root/reference through oracle/ref_shim.py) on seeded weights and
inputs.  Run in the build container only:  python oracle/make_golden.py

Each fixture stores the generator arguments (so tests can rebuild the exact inputs with
cs566-project-lightglue_b200/weights.py), a checksum of the weights, and the reference's outputs.
"""
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
OUT = os.path.join(ROOT, "tests", "golden")

CASES = {
    # name: (conf, batch, m, n, adaptive_fixture, image_size, correlated, full_la)
    "sp_b1_192": ({"filter_threshold": 0.0}, 1, 192, 192, False, (640, 480), 0.7, True),
    "sp_b2_160x224": ({"filter_threshold": 0.0}, 2, 160, 224, False, (640, 480), 0.7, True),
    "disk_b1_130x97": ({"filter_threshold": 0.0, "input_dim": 128}, 1, 130, 97, False, (1600, 1200), 0.7, True),
    "adaptive_both_512x400": ({"filter_threshold": 0.1, "depth_confidence": 0.95, "width_confidence": 0.95},
                              1, 512, 400, True, (640, 480), 0.7, True),
    "adaptive_width_300x400": ({"filter_threshold": 0.1, "width_confidence": 0.95}, 1, 300, 400, True, (640, 480), 0.7, True),
    "adaptive_depth_300x400": ({"filter_threshold": 0.0, "depth_confidence": 0.95}, 1, 300, 400, True, (640, 480), 0.7, True),
    "sp_b1_1024": ({"filter_threshold": 0.0}, 1, 1024, 1024, False, (1600, 1200), 0.7, False),
}


def weights_checksum(sd):
    return float(sum(v.double().abs().sum().item() for v in sd.values()))


def save(name, meta, arrays):
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), meta=json.dumps(meta), **arrays)


def forward_cases():
    for name, (conf, b, m, n, adapt, size, corr, full) in CASES.items():
        sd = W.make_state_dict(conf, seed=0, adaptive_fixture=adapt)
        model = ref_shim.build_reference(conf, sd)
        data = W.make_inputs(b, m, n, d_in=conf.get("input_dim", 256), seed=1, image_size=size, correlated=corr)
        pred = ref_shim.run_reference(model, data)
        la = pred["log_assignment"].numpy()
        arrays = {k: pred[k].numpy() for k in ("matches0", "matches1", "matching_scores0", "matching_scores1", "prune0", "prune1")}
        if full:
            arrays["log_assignment"] = la
        else:  # big case: strided sample + moments
            arrays["log_assignment_sample"] = la[:, ::17, ::19].copy()
            arrays["log_assignment_rowsum"] = la.astype(np.float64).sum(2)
        meta = dict(conf=conf, batch=b, m=m, n=n, adaptive_fixture=adapt, image_size=size, correlated=corr,
                    weight_seed=0, data_seed=1, weights_checksum=weights_checksum(sd),
                    la_shape=list(la.shape), source="gluefactory/models/matchers/lightglue.py:444-579 (reference run)")
        save("fwd_" + name, meta, arrays)
        print(name, la.shape, int((pred["matches0"] >= 0).sum()))


def varlen_case():
    """Config-4 style: drive the reference's own submodules with padding masks
    (lightglue.py:258-281) as SURVEY.md 8(c) prescribes."""
    conf = {"filter_threshold": 0.0, "input_dim": 128}
    b, npad = 3, 256
    len0, len1 = [256, 131, 77], [200, 256, 1]
    sd = W.make_state_dict(conf, seed=0)
    model = ref_shim.build_reference(conf, sd)
    lg = ref_shim.load_reference_module()
    data = W.make_inputs(b, npad, npad, d_in=128, seed=2, image_size=(1600, 1200), correlated=0.7)
    with torch.no_grad():
        k0 = lg.normalize_keypoints(data["keypoints0"], data["view0"]["image_size"])
        k1 = lg.normalize_keypoints(data["keypoints1"], data["view1"]["image_size"])
        d0, d1 = model.input_proj(data["descriptors0"]), model.input_proj(data["descriptors1"])
        cond = npad * torch.ones((b, 1))
        e0, e1 = model.posenc(k0, cond), model.posenc(k1, cond)
        ar = torch.arange(npad)
        mask0 = (ar[None] < torch.tensor(len0)[:, None])[:, None, :, None]
        mask1 = (ar[None] < torch.tensor(len1)[:, None])[:, None, :, None]
        for i in range(model.conf.n_layers):
            d0, d1 = model.transformers[i](d0, d1, e0, e1, mask0, mask1)
        arrays = {}
        for p in range(b):
            s, _ = model.log_assignment[-1](d0[p:p + 1, :len0[p]], d1[p:p + 1, :len1[p]])
            m0, m1, ms0, ms1 = lg.filter_matches(s, 0.0)
            arrays.update({f"log_assignment_{p}": s.numpy(), f"matches0_{p}": m0.numpy(), f"matches1_{p}": m1.numpy(),
                           f"matching_scores0_{p}": ms0.numpy(), f"matching_scores1_{p}": ms1.numpy()})
    meta = dict(conf=conf, batch=b, npad=npad, len0=len0, len1=len1, image_size=(1600, 1200), correlated=0.7,
                weight_seed=0, data_seed=2, weights_checksum=weights_checksum(sd),
                source="reference transformers[i].masked_forward + log_assignment[-1] + filter_matches")
    save("varlen_disk_b3_256", meta, arrays)
    print("varlen", [arrays[f"log_assignment_{p}"].shape for p in range(b)])


def filter_cases():
    """Known-answer vectors for filter_matches (lightglue.py:321-337) incl. ties / boundaries."""
    lg = ref_shim.load_reference_module()
    g = torch.Generator().manual_seed(7)
    cases = {}
    s = torch.randn(2, 9, 12, generator=g) - 2
    cases["random_small"] = (s, 0.1)
    s = torch.full((1, 6, 6), -1.0)
    cases["all_equal"] = (s, 0.2)
    s = torch.randn(1, 8, 8, generator=g)
    s[0, 2, :] = s[0, 2, 3]
    s[0, :, 5] = s[0, 1, 5]
    cases["ties"] = (s - 3, 0.0)
    s = torch.full((1, 5, 5), -5.0)
    for i in range(4):
        s[0, i, (i + 1) % 4] = float(np.log(0.25))
    cases["threshold_equal"] = (s, 0.25)
    s = torch.full((1, 4, 6), float("-inf"))
    s[0, 0, 1] = -0.5
    cases["neg_inf_rows"] = (s, 0.0)
    s = torch.randn(1, 2, 2, generator=g)
    cases["single_point"] = (s, 0.0)
    s = torch.randn(3, 70, 45, generator=g) * 3 - 4
    cases["rect_70x45"] = (s, 0.05)
    arrays, meta = {}, {}
    for name, (s, th) in cases.items():
        m0, m1, ms0, ms1 = lg.filter_matches(s, th)
        arrays.update({name + "_scores": s.numpy(), name + "_m0": m0.numpy(), name + "_m1": m1.numpy(),
                       name + "_ms0": ms0.numpy(), name + "_ms1": ms1.numpy()})
        meta[name] = th
    save("filter_matches_kat", dict(thresholds=meta, source="lightglue.py:321-337 (reference run)"), arrays)
    print("filter cases", list(cases))


if __name__ == "__main__":
    assert ref_shim.reference_available(), "needs /root/reference (build container only)"
    torch.set_num_threads(8)
    forward_cases()
    varlen_case()
    filter_cases()
