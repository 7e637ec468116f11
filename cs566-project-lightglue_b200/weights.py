"""Deterministic random-init weights with the reference's state_dict names and shapes.

There is no network for checkpoints, so parity tests, golden fixtures and the bench all
use seeded random-init weights.  The generator is independent of torch's nn.init call
order (each tensor is drawn from its own counter-based stream keyed by the parameter
name), so the reference module, the oracle and the CUDA path are all fed the very same
numbers through `load_state_dict` on any machine.

Names / shapes follow the reference (SURVEY.md 8(a) "Weights"):
  gluefactory/models/matchers/lightglue.py:57-61   posenc.Wr, posenc.condition_modulation
  lightglue.py:168-176                              self_attn.{Wqkv,out_proj,ffn.0,ffn.1,ffn.3}
  lightglue.py:203-211                              cross_attn.{to_qk,to_v,to_out,ffn.*}
  lightglue.py:303-304                              log_assignment.i.{matchability,final_proj}
  lightglue.py:99                                   token_confidence.j.token.0
  lightglue.py:370-371                              input_proj (only if input_dim != descriptor_dim)
Distributions mimic torch defaults: Linear U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight
and bias, LayerNorm (1, 0), posenc.Wr N(0, 1) (lightglue.py:58, gamma = 1).
"""
import zlib

import numpy as np
import torch


def _rng(seed: int, name: str) -> np.random.Generator:
    return np.random.Generator(np.random.Philox(key=[seed & 0xFFFFFFFF, zlib.crc32(name.encode())]))


def _uniform(seed, name, shape, bound):
    return torch.from_numpy(_rng(seed, name).uniform(-bound, bound, size=shape).astype(np.float32))


def _linear(sd, seed, prefix, out_f, in_f):
    b = 1.0 / np.sqrt(in_f)
    sd[prefix + ".weight"] = _uniform(seed, prefix + ".weight", (out_f, in_f), b)
    sd[prefix + ".bias"] = _uniform(seed, prefix + ".bias", (out_f,), b)


def make_state_dict(conf: dict = None, seed: int = 0, adaptive_fixture: bool = False) -> dict:
    """Build the full LightGlue state_dict (fp32, CPU).

    adaptive_fixture: sharpen the matchability / token-confidence heads so that the
    adaptive path actually prunes and exits at random init (SURVEY.md 8(c) recipe:
    matchability weight*4, bias=-1.5; token-confidence weight*4, bias=-1.0+0.9*i).
    """
    conf = dict(conf or {})
    d = int(conf.get("descriptor_dim", 256))
    L = int(conf.get("n_layers", 9))
    d_in = int(conf.get("input_dim", d))
    heads = int(conf.get("num_heads", 4))
    pe_in = 2 + 2 * int(bool(conf.get("add_scale_ori", False)))
    half = (d // heads) // 2
    sd = {}
    if d_in != d:
        _linear(sd, seed, "input_proj", d, d_in)
    sd["posenc.Wr.weight"] = torch.from_numpy(
        _rng(seed, "posenc.Wr.weight").standard_normal((half, pe_in)).astype(np.float32)
    )
    _linear(sd, seed, "posenc.condition_modulation", half, 1)
    for i in range(L):
        for blk in ("self_attn", "cross_attn"):
            p = f"transformers.{i}.{blk}"
            if blk == "self_attn":
                _linear(sd, seed, p + ".Wqkv", 3 * d, d)
                _linear(sd, seed, p + ".out_proj", d, d)
            else:
                _linear(sd, seed, p + ".to_qk", d, d)
                _linear(sd, seed, p + ".to_v", d, d)
                _linear(sd, seed, p + ".to_out", d, d)
            _linear(sd, seed, p + ".ffn.0", 2 * d, 2 * d)
            sd[p + ".ffn.1.weight"] = torch.ones(2 * d)
            sd[p + ".ffn.1.bias"] = torch.zeros(2 * d)
            _linear(sd, seed, p + ".ffn.3", d, 2 * d)
        _linear(sd, seed, f"log_assignment.{i}.matchability", 1, d)
        _linear(sd, seed, f"log_assignment.{i}.final_proj", d, d)
        if i < L - 1:
            _linear(sd, seed, f"token_confidence.{i}.token.0", 1, d)
    if adaptive_fixture:
        for i in range(L):
            sd[f"log_assignment.{i}.matchability.weight"] *= 4.0
            sd[f"log_assignment.{i}.matchability.bias"].fill_(-1.5)
            if i < L - 1:
                sd[f"token_confidence.{i}.token.0.weight"] *= 4.0
                sd[f"token_confidence.{i}.token.0.bias"].fill_(-1.0 + 0.9 * i)
    return sd


def make_inputs(batch: int, m: int, n: int, d_in: int = 256, seed: int = 1,
                image_size=(1600, 1200), correlated: float = 0.0) -> dict:
    """Seeded synthetic data dict in the reference's input contract (lightglue.py:363,448-456).

    keypoints U(0,w)xU(0,h) pixels, descriptors unit-norm Gaussian (SuperPoint-shaped,
    gluefactory_nonfree/superpoint.py:341-345), view{0,1}.image_size int64 (B,2).
    correlated>0 makes descriptors1 a noisy permutation of descriptors0 so that even a
    random-init network produces a non-trivial set of mutual matches.
    """
    g = np.random.Generator(np.random.Philox(key=[seed & 0xFFFFFFFF, 0xDA7A]))
    wh = np.asarray(image_size, dtype=np.float32)
    k0 = (g.random((batch, m, 2), dtype=np.float32) * wh).astype(np.float32)
    k1 = (g.random((batch, n, 2), dtype=np.float32) * wh).astype(np.float32)
    d0 = g.standard_normal((batch, m, d_in)).astype(np.float32)
    d1 = g.standard_normal((batch, n, d_in)).astype(np.float32)
    if correlated > 0:
        for b in range(batch):
            perm = g.permutation(max(m, n))
            src = perm[:n] % m
            d1[b] = correlated * d0[b, src] + (1 - correlated) * d1[b]
            k1[b] = np.clip(k0[b, src] + 8 * g.standard_normal((n, 2)).astype(np.float32), 0, wh - 1)
    d0 /= np.linalg.norm(d0, axis=-1, keepdims=True)
    d1 /= np.linalg.norm(d1, axis=-1, keepdims=True)
    size = np.tile(np.asarray(image_size, dtype=np.int64)[None], (batch, 1))
    return {
        "keypoints0": torch.from_numpy(k0),
        "keypoints1": torch.from_numpy(k1),
        "descriptors0": torch.from_numpy(d0),
        "descriptors1": torch.from_numpy(d1),
        "view0": {"image_size": torch.from_numpy(size.copy())},
        "view1": {"image_size": torch.from_numpy(size.copy())},
    }
