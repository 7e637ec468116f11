"""B200-native LightGlue matcher forward -- drop-in for the reference's
gluefactory/models/matchers/lightglue.py:340-579 (class LightGlue, eval-mode forward).

Same constructor (`LightGlue(conf)`, conf = dict or DictConfig merged over `default_conf`,
lightglue.py:341-361,369), same data dict in (`keypoints0/1`, `descriptors0/1`, optional
`view0/1.image_size`, lightglue.py:444-456), same pred dict out (lightglue.py:567-577), same
state_dict names (lightglue.py:380-396), selected through `__main_model__` like the reference
(lightglue.py:666; gluefactory/models/__init__.py:20-25):  `model.matcher.name: b200glue.matcher`.

The hot subsystems run as hand-written sm_100a kernels behind the C ABI in include/lg_b200.h
(positional encoding, split-bf16 tcgen05 linears with the rotary / head split / residual fused
into their epilogues, LayerNorm + GELU, tcgen05 flash attention, TF32x3 assignment, match filter,
confidence / pruning); see engine.py for the per-layer launch sequence.  There is no CPU path
and no fallback: inputs must live on a CUDA device and the shared library must be present.
"""
import math
from pathlib import Path

import numpy as np
import torch
from torch import nn

from . import _cabi as C


class Conf(dict):
    """Attribute-access dict standing in for an OmegaConf DictConfig (omegaconf is optional)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def _to_plain(c):
    if c is None:
        return {}
    if isinstance(c, dict):
        return {k: _to_plain(v) if isinstance(v, dict) else v for k, v in c.items()}
    try:  # DictConfig
        from omegaconf import OmegaConf

        return OmegaConf.to_container(c, resolve=True)
    except ImportError:
        return dict(c)


def merge_conf(default: dict, user) -> Conf:
    out = Conf()
    user = _to_plain(user)
    for k, v in default.items():
        out[k] = merge_conf(v, user.get(k)) if isinstance(v, dict) else user.get(k, v)
    for k, v in user.items():  # unknown keys are tolerated (no struct mode in the reference)
        if k not in out:
            out[k] = Conf(v) if isinstance(v, dict) else v
    return out


class _PosEnc(nn.Module):  # parameter holder: lightglue.py:50-61
    def __init__(self, m_in, half):
        super().__init__()
        self.Wr = nn.Linear(m_in, half, bias=False)
        nn.init.normal_(self.Wr.weight.data, mean=0, std=1.0)
        self.condition_modulation = nn.Linear(1, half)


def _ffn(d):
    return nn.Sequential(nn.Linear(2 * d, 2 * d), nn.LayerNorm(2 * d, elementwise_affine=True), nn.GELU(),
                         nn.Linear(2 * d, d))


class _SelfBlock(nn.Module):  # lightglue.py:159-176
    def __init__(self, d):
        super().__init__()
        self.Wqkv = nn.Linear(d, 3 * d)
        self.out_proj = nn.Linear(d, d)
        self.ffn = _ffn(d)


class _CrossBlock(nn.Module):  # lightglue.py:194-211
    def __init__(self, d):
        super().__init__()
        self.to_qk = nn.Linear(d, d)
        self.to_v = nn.Linear(d, d)
        self.to_out = nn.Linear(d, d)
        self.ffn = _ffn(d)


class _Layer(nn.Module):  # lightglue.py:252-256
    def __init__(self, d):
        super().__init__()
        self.self_attn = _SelfBlock(d)
        self.cross_attn = _CrossBlock(d)


class _Assignment(nn.Module):  # lightglue.py:299-304
    def __init__(self, d):
        super().__init__()
        self.matchability = nn.Linear(d, 1)
        self.final_proj = nn.Linear(d, d)


class _TokenConfidence(nn.Module):  # lightglue.py:96-99
    def __init__(self, d):
        super().__init__()
        self.token = nn.Sequential(nn.Linear(d, 1), nn.Sigmoid())


class LightGlue(nn.Module):
    default_conf = {
        "name": "lightglue",
        "input_dim": 256,
        "add_scale_ori": False,
        "descriptor_dim": 256,
        "n_layers": 9,
        "num_heads": 4,
        "flash": False,
        "mp": False,
        "depth_confidence": -1,
        "width_confidence": -1,
        "filter_threshold": 0.0,
        "checkpointed": False,
        "weights": None,
        "weights_from_version": "v0.1_arxiv",
        "loss": {"gamma": 1.0, "fn": "nll", "nll_balancing": 0.5},
        # namespaced additions (existing YAMLs stay valid)
        # condition: "padded" = the reference's own behaviour on a padded batch (every image conditioned on the padded
        # length, lightglue.py:490-491); "valid" = every image of a batch with num_keypoints0/1 conditioned on ITS OWN
        # count, so that the batch reproduces the reference's one-pair-per-call results (pipeline.pack_pairs)
        "b200": {"precision": "bf16", "cuda_graph": True, "engine_cache": 2, "condition": "padded"},
    }
    required_data_keys = ["keypoints0", "keypoints1", "descriptors0", "descriptors1"]

    def __init__(self, conf=None) -> None:
        super().__init__()
        self.conf = conf = merge_conf(self.default_conf, conf)
        d, h, L = conf.descriptor_dim, conf.num_heads, conf.n_layers
        if d != 256 or h != 4:
            raise ValueError("the sm_100a kernels are built for descriptor_dim=256, num_heads=4 (no fallback)")
        if conf.b200.precision not in ("bf16", "fp32"):
            raise ValueError(f"b200.precision={conf.b200.precision!r}: 'bf16' (tensor-core throughput path) or 'fp32'")
        if conf.b200.condition not in ("padded", "valid"):
            raise ValueError(f"b200.condition={conf.b200.condition!r}: 'padded' (reference semantics) or 'valid'")
        self.input_proj = nn.Linear(conf.input_dim, d, bias=True) if conf.input_dim != d else nn.Identity()
        self.posenc = _PosEnc(2 + 2 * int(bool(conf.add_scale_ori)), (d // h) // 2)
        self.transformers = nn.ModuleList([_Layer(d) for _ in range(L)])
        self.log_assignment = nn.ModuleList([_Assignment(d) for _ in range(L)])
        self.token_confidence = nn.ModuleList([_TokenConfidence(d) for _ in range(L - 1)])
        # the attribute the fork forgot: thresholds as an fp32 table
        self.register_buffer(
            "confidence_thresholds",
            torch.tensor([self.confidence_threshold(i) for i in range(L)], dtype=torch.float32), persistent=False)
        if conf.weights is not None:
            self._load_weights(conf.weights)
        self.eval()

    # ------------------------------------------------------------------ weights (lightglue.py:400-430)
    def _load_weights(self, weights):
        path = Path(weights)
        if not path.exists():
            raise FileNotFoundError(f"weights {weights!r}: only local files are supported (no network on the box)")
        state = torch.load(str(path), map_location="cpu")
        if "model" in state and isinstance(state["model"], dict):
            state = state["model"]
        self.load_state_dict(self.rename_legacy_keys(state, self.conf.n_layers), strict=False)

    @staticmethod
    def rename_legacy_keys(state, n_layers):
        for i in range(n_layers):  # lightglue.py:425-429
            state = {k.replace(f"self_attn.{i}", f"transformers.{i}.self_attn"): v for k, v in state.items()}
            state = {k.replace(f"cross_attn.{i}", f"transformers.{i}.cross_attn"): v for k, v in state.items()}
        return state

    def confidence_threshold(self, layer_index: int) -> float:  # lightglue.py:581-584
        return float(np.clip(0.8 + 0.1 * np.exp(-4.0 * layer_index / self.conf.n_layers), 0, 1))

    # ------------------------------------------------------------------ packed weights / engine
    def _packed(self):
        from .engine import PackedWeights

        pk = getattr(self, "_pk", None)
        if pk is None or pk.signature != PackedWeights.sign(self):
            pk = self._pk = PackedWeights(self)
        return pk

    def _engine(self, B, M, N, device, cached=False):
        """Activation buffers (+ captured CUDA graph) for one problem shape; the batched path keeps the last few."""
        from .engine import TileEngine

        pk = self._packed()
        if not cached:
            return TileEngine(pk, B, M, N, device, self.conf.num_heads)
        cache = self.__dict__.setdefault("_engines", {})
        key = (B, M, N, str(device))
        eng = cache.pop(key, None)
        if eng is None or eng.pk is not pk:
            eng = TileEngine(pk, B, M, N, device, self.conf.num_heads)
        cache[key] = eng  # most recently used last
        while len(cache) > max(int(self.conf.b200.engine_cache), 1):
            cache.pop(next(iter(cache)))
        return eng

    def _assign(self, i, eng, d0, d1, len0, len1, th=None):
        """MatchAssignment.forward (lightglue.py:306-315): final_proj on tcgen05 straight into split tiles, the
        matchability logits, then the fused similarity + double log-softmax kernel on those tiles.  With `th` the
        same launch also gathers the row / column argmax keys and filter_matches (lightglue.py:321-337) is finished
        from them: returns (scores, matches)."""
        head = self.log_assignment[i]
        a_tiles, b_tiles = eng.final_proj_tiles(i)
        wm = head.matchability.weight.reshape(-1)
        z0 = C.logit_rows(d0, wm, head.matchability.bias)
        z1 = C.logit_rows(d1, wm, head.matchability.bias)
        scale = float(self.conf.descriptor_dim) ** -0.5  # d^-1/4 on both operands (lightglue.py:311-312)
        return C.assign_tiles(a_tiles, b_tiles, eng.tp0, eng.tp1, z0, z1, th, len0, len1, scale=scale)

    def _forward_fp32(self, desc0, desc1, cs0, cs1, len0, len1):
        """The accuracy mode (engine_fp32.py): IEEE fp32 layer stack, TF32x3 similarity; 1e-4 bound on log_assignment."""
        from .engine import PackedWeights
        from .engine_fp32 import Fp32Engine, Fp32Weights, forward_adaptive_fp32, max_pairs

        w = getattr(self, "_w32", None)
        if w is None or w.signature != PackedWeights.sign(self):
            w = self._w32 = Fp32Weights(self)
            w.signature = PackedWeights.sign(self)
        b, m, n = desc0.shape[0], desc0.shape[1], desc1.shape[1]
        L, dev = self.conf.n_layers, desc0.device
        if self.conf.depth_confidence > 0 or self.conf.width_confidence > 0:
            return self._forward_fp32_adaptive(w, desc0, desc1, cs0, cs1, len0, len1)
        step = min(b, max_pairs(m, n))
        outs = []
        for lo in range(0, b, step):
            hi = min(b, lo + step)
            cut = lambda t: None if t is None else t[lo:hi].contiguous()
            # the buffers of the last two shapes are kept (a batch cut into chunks ends with a shorter one; the
            # materialised scores alone are up to ~1 GB)
            cache = self.__dict__.setdefault("_engines32", {})
            key = (hi - lo, m, n, str(dev))
            eng = cache.pop(key, None)
            if eng is None or eng.w is not w:
                while len(cache) > 1:
                    cache.pop(next(iter(cache)))
                eng = Fp32Engine(w, hi - lo, m, n, dev)
            cache[key] = eng
            l0, l1 = cut(len0), cut(len1)
            d0, d1 = eng.forward(cut(desc0), cut(desc1), (cut(cs0[0]), cut(cs0[1])), (cut(cs1[0]), cut(cs1[1])), l0, l1, L)
            scores, (m0, m1, ms0, ms1) = eng.assign(L - 1, self.conf.filter_threshold, l0, l1)
            outs.append((m0, m1, ms0, ms1, d0.clone()[:, None], d1.clone()[:, None], scores))
        m0, m1, ms0, ms1, r0, r1, scores = (torch.cat(t, 0) if len(outs) > 1 else t[0] for t in zip(*outs))
        return {
            "matches0": m0, "matches1": m1, "matching_scores0": ms0, "matching_scores1": ms1,
            "ref_descriptors0": r0, "ref_descriptors1": r1, "log_assignment": scores,
            "prune0": torch.ones_like(ms0) * L, "prune1": torch.ones_like(ms1) * L,
        }

    def _forward_fp32_adaptive(self, w, desc0, desc1, cs0, cs1, len0, len1):
        """depth / width confidence in the fp32 mode: the reference's one-pair procedure (lightglue.py:502-565), pair by
        pair; a pair of a padded batch runs on its valid points.  B == 1 returns the reference's contract, B > 1 the
        batched one (exit_layer (B,), log_assignment a list of per-pair (1, M'+1, N'+1) matrices)."""
        from .engine_fp32 import forward_adaptive_fp32

        b, m, n = desc0.shape[0], desc0.shape[1], desc1.shape[1]
        l0 = [m] * b if len0 is None else [int(v) for v in len0.tolist()]
        l1 = [n] * b if len1 is None else [int(v) for v in len1.tolist()]
        preds = []
        for p in range(b):
            a, c = l0[p], l1[p]
            cut = lambda t, k: t[p:p + 1, :k].contiguous()
            preds.append(forward_adaptive_fp32(self, w, cut(desc0, a), cut(desc1, c), (cut(cs0[0], a), cut(cs0[1], a)),
                                               (cut(cs1[0], c), cut(cs1[1], c))))
        if b == 1 and len0 is None:
            return preds[0]
        out = {"log_assignment": [q["log_assignment"] for q in preds],
               "exit_layer": torch.tensor([q["exit_layer"] for q in preds], dtype=torch.int64)}
        for key, width, fill in (("matches0", m, -1), ("matches1", n, -1), ("matching_scores0", m, 0), ("matching_scores1", n, 0),
                                 ("prune0", m, 0), ("prune1", n, 0)):
            t = preds[0][key].new_full((b, width), fill)
            for p, q in enumerate(preds):
                t[p, : q[key].shape[1]] = q[key][0]
            out[key] = t
        out["ref_descriptors0"] = desc0.new_zeros((b, 0, m, self.conf.descriptor_dim))
        out["ref_descriptors1"] = desc1.new_zeros((b, 0, n, self.conf.descriptor_dim))
        return out

    @staticmethod
    def _image_size(data, view, like):
        v = data.get(view)
        size = None if v is None else v.get("image_size")
        if size is None:
            return None
        if not isinstance(size, torch.Tensor):
            size = torch.tensor(size)
        size = size.to(device=like.device, dtype=torch.float32)
        if size.dim() == 1:
            size = size[None].expand(like.shape[0], 2)
        return size.contiguous()

    @staticmethod
    def _lens(data, key, n, device):
        v = data.get(key)
        if v is None:
            return None
        return torch.as_tensor(v, device=device).to(torch.int32).clamp(0, n).contiguous()

    # ------------------------------------------------------------------ forward (lightglue.py:444-579)
    @torch.no_grad()
    def forward(self, data: dict) -> dict:
        for key in self.required_data_keys:
            assert key in data, f"Missing key {key} in data"
        kpts0 = data["keypoints0"]
        if not kpts0.is_cuda:
            raise RuntimeError("b200glue LightGlue runs on CUDA (sm_100a) only; move the batch to the GPU")
        with torch.cuda.device(kpts0.device), C.pinned_stream():  # the C ABI launches on the current device's current stream
            return self._forward(data)

    def _forward(self, data: dict) -> dict:
        conf = self.conf
        kpts0, kpts1 = data["keypoints0"], data["keypoints1"]
        b, m, _ = kpts0.shape
        _, n, _ = kpts1.shape
        dev = kpts0.device
        desc0 = data["descriptors0"].contiguous().float()
        desc1 = data["descriptors1"].contiguous().float()
        assert desc0.shape[-1] == conf.input_dim and desc1.shape[-1] == conf.input_dim
        len0 = self._lens(data, "num_keypoints0", m, dev)
        len1 = self._lens(data, "num_keypoints1", n, dev)
        if (len0 is None) != (len1 is None):  # one side given: the other is full length
            len0 = torch.full((b,), m, device=dev, dtype=torch.int32) if len0 is None else len0
            len1 = torch.full((b,), n, device=dev, dtype=torch.int32) if len1 is None else len1
        pe = self.posenc
        wc = pe.condition_modulation.weight.reshape(-1)
        so0 = so1 = (None, None)
        if conf.add_scale_ori:  # 4-d positional input [x, y, scale, ori] (lightglue.py:458-476)
            flat = lambda t, k: t.reshape(b, k).float().contiguous()
            so0 = (flat(data["scales0"], m), flat(data["oris0"], m))
            so1 = (flat(data["scales1"], n), flat(data["oris1"], n))
        per_image = conf.b200.condition == "valid" and len0 is not None
        cond0 = len0.float() if per_image else float(m)
        cond1 = len1.float() if per_image else float(n)
        cs0 = C.posenc(kpts0.float().contiguous(), self._image_size(data, "view0", kpts0), pe.Wr.weight, wc,
                       pe.condition_modulation.bias, cond0, *so0, lens=len0 if per_image else None)
        cs1 = C.posenc(kpts1.float().contiguous(), self._image_size(data, "view1", kpts1), pe.Wr.weight, wc,
                       pe.condition_modulation.bias, cond1, *so1, lens=len1 if per_image else None)
        do_early_stop = conf.depth_confidence > 0
        do_point_pruning = conf.width_confidence > 0
        if conf.b200.precision == "fp32":
            return self._forward_fp32(desc0, desc1, cs0, cs1, len0, len1)
        eng = self._engine(b, m, n, dev, cached=True)  # (the adaptive path moves to smaller, uncached engines as it prunes)
        eng.load(desc0, desc1, cs0, cs1)

        if do_early_stop or do_point_pruning:
            from .adaptive import forward_adaptive_batched, forward_adaptive_one

            if b == 1 and len0 is None:
                return forward_adaptive_one(self, eng, cs0, cs1)
            return forward_adaptive_batched(self, eng, cs0, cs1, len0, len1)

        L = conf.n_layers
        eng.run_layers(L, len0, len1, use_graph=bool(conf.b200.cuda_graph))
        desc0, desc1 = eng.rows()
        scores, (m0, m1, ms0, ms1) = self._assign(L - 1, eng, desc0, desc1, len0, len1, th=conf.filter_threshold)
        return {
            "matches0": m0, "matches1": m1,
            "matching_scores0": ms0, "matching_scores1": ms1,
            "ref_descriptors0": desc0[:, None], "ref_descriptors1": desc1[:, None],
            "log_assignment": scores,
            "prune0": torch.ones_like(ms0) * L, "prune1": torch.ones_like(ms1) * L,
        }


__main_model__ = LightGlue
