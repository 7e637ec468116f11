"""TEST INFRASTRUCTURE.  Goldens at the sizes of BASELINE configs 3 and 4, from the UNMODIFIED reference module
(imported from /root/reference through oracle/ref_shim.py).  Run in the build container only:
    python oracle/make_golden_scale.py

  fwd_adaptive_both_2048.npz   config 3: one pair of 2048 x 2048 keypoints, depth = width confidence 0.95.  Besides the
      outputs, the reference's OWN per-layer score tensors are recorded through hooks on TokenConfidence.forward
      (lightglue.py:102-107), MatchAssignment.get_matchability (:317-318) and LightGlue.get_pruning_mask (:586-593):
      token confidences, matchability scores, keep masks and the surviving original indices ind0 / ind1 (:536-547).
      The kernel tests feed exactly these tensors to lg_prune_plan (pruning indices bit-exact "when fed the reference's
      own score tensors") and use the keep margins to tell a legitimate flip of a marginal token from an error.
  trace_adaptive_*.npz         the same per-layer tensors for the three small adaptive fixtures of make_golden.py.
  varlen_disk_b2_4096.npz      config 4: DISK-shaped 128-d descriptors, two pairs padded to 4096 keypoints with unequal
      valid counts, through the reference's own masked submodules (:258-281) like the small varlen golden; the
      (M_b+1) x (N_b+1) matrices are too big for a fixture: a strided sample, row sums, and the full match vectors.
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
from oracle.make_golden import save, weights_checksum  # noqa: E402

W = importlib.import_module("cs566-project-lightglue_b200.weights")


def adaptive_case(name, conf, m, n, size, trace_only=False):
    """trace_only: the outputs already live in tests/golden/fwd_<name>.npz (oracle/make_golden.py); write only the
    per-layer score tensors to trace_<name>.npz after checking that this run reproduces that fixture."""
    sd = W.make_state_dict(conf, seed=0, adaptive_fixture=True)
    model = ref_shim.build_reference(conf, sd)
    data = W.make_inputs(1, m, n, seed=1, image_size=size, correlated=0.7)
    rec = {"conf": [], "match": [], "keep": []}
    for i, tc in enumerate(model.token_confidence):
        orig = tc.forward

        def fwd(d0, d1, _orig=orig):
            t0, t1 = _orig(d0, d1)
            rec["conf"].append((t0.clone(), t1.clone()))
            return t0, t1

        tc.forward = fwd
    for i, head in enumerate(model.log_assignment):
        orig = head.get_matchability

        def gm(desc, _orig=orig):
            s = _orig(desc)
            rec["match"].append(s.clone())
            return s

        head.get_matchability = gm
    orig_mask = model.get_pruning_mask

    def mask(confidences, scores, layer_index):
        k = orig_mask(confidences, scores, layer_index)
        rec["keep"].append((layer_index, k.clone()))
        return k

    model.get_pruning_mask = mask
    orig_stop = model.check_if_stop
    stops = []

    def stop(c0, c1, layer_index, num_points):
        r = orig_stop(c0, c1, layer_index, num_points)
        stops.append((layer_index, bool(r)))
        return r

    model.check_if_stop = stop
    pred = ref_shim.run_reference(model, data)
    la = pred["log_assignment"].numpy()
    arrays = {k: pred[k].numpy() for k in ("matches0", "matches1", "matching_scores0", "matching_scores1", "prune0", "prune1")}
    arrays["log_assignment_sample"] = la[:, ::17, ::19].copy()
    arrays["log_assignment_rowsum"] = la.astype(np.float64).sum(2)
    n_conf, n_prune = len(rec["conf"]), len(rec["keep"]) // 2
    thr = [float(model.confidence_thresholds[i]) for i in range(model.conf.n_layers)]
    for i, (t0, t1) in enumerate(rec["conf"]):
        arrays[f"L{i}_conf0"], arrays[f"L{i}_conf1"] = t0.numpy(), t1.numpy()
    for i in range(n_prune):
        (l0, k0), (l1, k1) = rec["keep"][2 * i], rec["keep"][2 * i + 1]
        assert l0 == l1 == i
        arrays[f"L{i}_match0"], arrays[f"L{i}_match1"] = rec["match"][2 * i].numpy(), rec["match"][2 * i + 1].numpy()
        arrays[f"L{i}_keep0"], arrays[f"L{i}_keep1"] = k0.numpy(), k1.numpy()
    if trace_only:
        old = np.load(os.path.join(ROOT, "tests", "golden", "fwd_" + name + ".npz"))
        for k in ("matches0", "prune0", "prune1", "log_assignment"):
            assert np.array_equal(old[k], pred[k].numpy()), k
        arrays = {k: v for k, v in arrays.items() if k.startswith("L")}
    meta = dict(conf=conf, batch=1, m=m, n=n, adaptive_fixture=True, image_size=list(size), correlated=0.7,
                weight_seed=0, data_seed=1, weights_checksum=weights_checksum(sd), la_shape=list(la.shape),
                layers_scored=n_conf, layers_pruned=n_prune, confidence_thresholds=thr,
                exit_layer=next((i for i, r in stops if r), model.conf.n_layers - 1), stop_decisions=stops,
                source="gluefactory/models/matchers/lightglue.py:444-579 (reference run, hooks on :102-107, :317-318, :586-593)")
    save(("trace_" if trace_only else "fwd_") + name, meta, arrays)
    print(name, "la", la.shape, "scored layers", n_conf, "pruned layers", n_prune, "matches",
          int((pred["matches0"] >= 0).sum()), "exit", meta["exit_layer"])


def varlen_4096():
    conf = {"filter_threshold": 0.0, "input_dim": 128}
    b, npad = 2, 4096
    len0, len1 = [4096, 2311], [2950, 4096]
    sd = W.make_state_dict(conf, seed=0)
    model = ref_shim.build_reference(conf, sd)
    lg = ref_shim.load_reference_module()
    data = W.make_inputs(b, npad, npad, d_in=128, seed=3, image_size=(1600, 1200), correlated=0.7)
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
            print("  layer", i, flush=True)
        arrays = {}
        for p in range(b):
            s, _ = model.log_assignment[-1](d0[p:p + 1, :len0[p]], d1[p:p + 1, :len1[p]])
            m0, m1, ms0, ms1 = lg.filter_matches(s, 0.0)
            la = s.numpy()
            arrays.update({f"log_assignment_sample_{p}": la[:, ::31, ::37].copy(),
                           f"log_assignment_rowsum_{p}": la.astype(np.float64).sum(2),
                           f"log_assignment_last_{p}": la[:, -3:, :].copy(),  # incl. the dustbin row
                           f"matches0_{p}": m0.numpy().astype(np.int32), f"matches1_{p}": m1.numpy().astype(np.int32),
                           f"matching_scores0_{p}": ms0.numpy(), f"matching_scores1_{p}": ms1.numpy()})
            print("  pair", p, la.shape, int((m0 >= 0).sum()), "matches", flush=True)
    meta = dict(conf=conf, batch=b, npad=npad, len0=len0, len1=len1, image_size=(1600, 1200), correlated=0.7,
                weight_seed=0, data_seed=3, weights_checksum=weights_checksum(sd), sample_stride=[31, 37],
                source="reference transformers[i].masked_forward + log_assignment[-1] + filter_matches")
    save("varlen_disk_b2_4096", meta, arrays)


if __name__ == "__main__":
    assert ref_shim.reference_available(), "needs /root/reference (build container only)"
    torch.set_num_threads(max(1, (os.cpu_count() or 8)))
    which = sys.argv[1:] or ["adaptive", "traces", "varlen"]
    if "adaptive" in which:
        adaptive_case("adaptive_both_2048", {"filter_threshold": 0.0, "depth_confidence": 0.95, "width_confidence": 0.95},
                      2048, 2048, (1600, 1200))
    if "traces" in which:  # per-layer tensors behind the three small adaptive fixtures of oracle/make_golden.py
        from oracle.make_golden import CASES
        for name, (conf, b, m, n, adapt, size, corr, full) in CASES.items():
            if adapt:
                adaptive_case(name, conf, m, n, size, trace_only=True)
    if "varlen" in which:
        varlen_4096()
