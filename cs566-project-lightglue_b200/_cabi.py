"""ctypes binding of liblg_b200.so (C ABI declared in include/lg_b200.h).

PyTorch is only plumbing here: it owns device memory and the stream; every wrapper passes raw
device pointers + sizes + the current CUDA stream handle.  There is NO fallback: if the shared
library is missing or a call returns non-zero, a RuntimeError is raised.
"""
import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LG_B200_LIB") or os.path.join(_HERE, "liblg_b200.so")  # env override: kernel experiments
TILE = 128
TILE_BYTES = 16384

_c = ctypes
_P, _I, _F, _L = _c.c_void_p, _c.c_int, _c.c_float, _c.c_int64

# name -> (restype, argtypes); must list every symbol include/lg_b200.h declares
SIGNATURES = {
    "lg_version": (_I, []),
    "lg_source_id": (_c.c_char_p, []),
    "lg_posenc": (_I, [_P, _P, _P, _P, _P, _F, _P, _P, _I, _I, _P]),
    "lg_posenc_scale_ori": (_I, [_P, _P, _P, _P, _P, _P, _P, _F, _P, _P, _I, _I, _P]),
    "lg_keypoint_extent": (_I, [_P, _P, _I, _I, _P]),
    "lg_keypoint_extent_valid": (_I, [_P, _P, _P, _I, _I, _P]),
    "lg_posenc_per_image": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P]),
    "lg_pack_self_qkv": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "lg_pack_cross_qkv": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "lg_attention": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _P]),
    "lg_assign_ws_bytes": (_c.c_size_t, [_I, _I, _I]),
    "lg_assign": (_I, [_P, _P, _P, _P, _P, _P, _F, _P, _P, _I, _I, _I, _P]),
    "lg_token_scores": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P]),
    "lg_prune_plan": (_I, [_P, _P, _P, _F, _F, _P, _P, _I, _I, _P]),
    "lg_prune_gather": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "lg_score_prune_tiles": (_I, [_P] * 21 + [_F, _F, _I, _P] + [_I] * 7 + [_P]),
    "lg_score_prune_ws_bytes": (_c.c_size_t, [_I, _I, _I]),
    "lg_filter_ws_bytes": (_c.c_size_t, [_I, _I, _I]),
    "lg_pack_weight": (_I, [_P, _P, _I, _I, _I, _P]),
    "lg_rows_to_tiles": (_I, [_P, _P, _I, _I, _I, _I, _P]),
    "lg_tiles_to_rows": (_I, [_P, _P, _I, _I, _I, _I, _P]),
    "lg_rotary_tiles": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "lg_ln_gelu_tiles": (_I, [_P, _P, _P, _F, _P, _I, _P]),
    "lg_ffn_ln_gelu": (_I, [_P, _I, _P, _I, _P, _P, _P, _P, _F, _P, _I, _P]),
    "lg_linear_split": (_I, [_P, _I, _P, _I, _I, _P, _P, _P, _P, _I, _I, _P]),
    "lg_linear_f32": (_I, [_P, _I, _P, _I, _I, _P, _P, _P, _L, _I, _I, _P]),
    "lg_linear_f16": (_I, [_P, _I, _P, _I, _I, _P, _P, _P, _I, _I, _P]),
    "lg_linear_qkv_self": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "lg_linear_qkv_cross": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "lg_attention_tiles": (_I, [_P, _P, _P, _P, _I, _P, _P, _I, _I, _I, _I, _I, _I, _F, _P]),
    "lg_attention_tiles2": (_I, [_P, _P, _P, _I, _P, _P, _I, _I, _I, _I, _P, _P, _P, _I, _P, _P, _I, _I, _I, _I, _P, _I, _I,
                                 _F, _P]),
    "lg_filter_matches": (_I, [_P, _L, _P, _P, _F, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "lg_filter_matches_static": (_I, [_P, _L, _P, _P, _F, _P, _P, _P, _P, _I, _I, _I, _P]),
    "lg_assign_keys": (_I, [_P, _P, _P, _P, _P, _P, _F, _P, _P, _P, _I, _I, _I, _P]),
    "lg_filter_from_keys": (_I, [_P, _P, _P, _F, _P, _P, _P, _P, _I, _I, _I, _P]),
    "lg_assign_tiles_ws_bytes": (_c.c_size_t, [_I, _I, _I]),
    "lg_assign_tiles": (_I, [_P, _P, _I, _I, _P, _P, _P, _P, _F, _P, _L, _P, _P, _I, _I, _I, _P]),
    "lg_logit_rows": (_I, [_P, _P, _P, _P, _L, _P]),
    "lg_nn_assign": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "lg_nn_ws_bytes": (_c.c_size_t, [_I, _I, _I]),
    "lg_nn_match": (_I, [_P, _P, _P, _F, _F, _I, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "lg_gemm_f32": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _L, _L, _L, _L, _I, _F, _I, _I, _P, _P, _P, _P]),
    "lg_softmax_rows_f32": (_I, [_P, _P, _P, _I, _I, _I, _I, _L, _P]),
    "lg_rotary_rows_f32": (_I, [_P, _P, _P, _L, _L, _I, _I, _I, _P]),
    "lg_ln_gelu_rows_f32": (_I, [_P, _P, _P, _F, _L, _I, _P]),
}

_lib = None

# kernels launched by this library since import (bench.py reports it as gpu_launches)
LAUNCHES = 0
# when a list, attention() appends (start_event, end_event, flops) around every launch
ATTN_EVENTS = None


def _count(n):
    global LAUNCHES
    LAUNCHES += n


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into liblg_b200.so (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc"), "-j8"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building liblg_b200.so failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout)
    return LIB_PATH


def source_id() -> str:
    """What lg_source_id() of a library built from THIS tree returns (the Makefile computes the same digest)."""
    import glob
    import hashlib
    csrc = os.path.join(_HERE, "csrc")
    files = sorted(glob.glob(os.path.join(csrc, "lg_*.cu"))) + sorted(glob.glob(os.path.join(csrc, "lg_*.cuh")))
    files.append(os.path.join(os.path.dirname(_HERE), "include", "lg_b200.h"))
    h = hashlib.sha256()
    for f in files:
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def _built_id():
    """Digest embedded in the library file (read without loading it), or None."""
    import re
    with open(LIB_PATH, "rb") as fh:
        m = re.search(rb"LG_SOURCE_ID=([0-9a-f]{16})", fh.read())
    return None if m is None else m.group(1).decode()


def lib():
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: the CUDA extension is required (no CPU / PyTorch fallback). "
                "Run `python -c 'import __graft_entry__ as g; g.build()'` or `make -C cs566-project-lightglue_b200/csrc`."
            )
        check = not os.environ.get("LG_B200_LIB")  # (kernel experiments pick their own build)
        if check and _built_id() != source_id():   # a library older than csrc/: rebuild once, never load it silently
            try:
                build()
            except Exception as e:
                raise RuntimeError(f"{LIB_PATH} is stale (built from sources {_built_id()}, the tree is {source_id()}) "
                                   f"and rebuilding failed: {e}") from e
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype, fn.argtypes = res, args
        if check and l.lg_source_id().decode() != source_id():
            raise RuntimeError(f"{LIB_PATH} is stale: built from sources {l.lg_source_id().decode()}, the tree is "
                               f"{source_id()}. Rebuild: `make -C cs566-project-lightglue_b200/csrc`.")
        _lib = l
    return _lib


def _check(code: int, what: str):
    if code != 0:
        kind = {-1: "null pointer", -2: "bad shape", -3: "misaligned pointer",
                -4: "problem larger than the fixed capacity of this non-allocating entry point",
                719: "cudaError 719 (launch failure: a device trap -- the kernels abort through mbar_wait's spin limit "
                     "instead of hanging when a pipeline stalls -- or a fault of an EARLIER asynchronous launch)",
                }.get(code, f"cudaError {code}")
        raise RuntimeError(f"{what} failed: {kind}")


def _ptr(t):
    return None if t is None else t.data_ptr()


_STREAM = None  # stream handle pinned for the duration of one forward (see pinned_stream)


def _stream():
    """Current stream of the CURRENT device: callers run under torch.cuda.device(<device of their tensors>)
    (matcher.forward, HostPipeline.run), so that a model on cuda:1 launches on cuda:1."""
    if _STREAM is not None:
        return _STREAM
    return torch.cuda.current_stream().cuda_stream


class pinned_stream:
    """`with pinned_stream():` -- look the current stream up ONCE for the ~100 launches of a forward
    (torch.cuda.current_stream() costs ~15 us: 1 ms of host time per one-pair call).  `pinned_stream(None)` suspends
    the pin (CUDA graph capture runs on its own stream)."""

    def __init__(self, handle="current"):
        self.handle = handle

    def __enter__(self):
        global _STREAM
        self.prev = _STREAM
        _STREAM = torch.cuda.current_stream().cuda_stream if self.handle == "current" else self.handle

    def __exit__(self, *exc):
        global _STREAM
        _STREAM = self.prev


def _f32c(t, name):
    assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous(), f"{name}: want contiguous cuda float32"
    return t


def _i32(t):
    if t is None:
        return None
    assert t.is_cuda and t.dtype == torch.int32 and t.is_contiguous()
    return t


def tiles_pad(n: int) -> int:
    t = (n + TILE - 1) // TILE
    return t + (t & 1)


def version() -> int:
    return lib().lg_version()


def posenc(kpts, size, wr, wc, bc, cond, scales=None, oris=None, lens=None):
    """-> (cos, sin) compact tables (B,N,32). size (B,2) f32 or None (extent of the points, of the first lens[b] points
    when `lens` is given); scales / oris (B,N): the 4-d positional input of add_scale_ori (wr is then (32,4)).
    cond: the keypoint count as a float (the reference's scalar condition), or a (B,) tensor = one count per image."""
    B, N, _ = kpts.shape
    _f32c(kpts, "kpts")
    if size is None:
        size = torch.empty(B, 2, device=kpts.device, dtype=torch.float32)
        if lens is None:
            _check(lib().lg_keypoint_extent(_ptr(kpts), _ptr(size), B, N, _stream()), "lg_keypoint_extent")
        else:
            _check(lib().lg_keypoint_extent_valid(_ptr(kpts), _ptr(_i32(lens)), _ptr(size), B, N, _stream()),
                   "lg_keypoint_extent_valid")
        _count(1)
    cos = torch.empty(B, N, 32, device=kpts.device, dtype=torch.float32)
    sin = torch.empty_like(cos)
    assert wr.shape == ((32, 4) if scales is not None else (32, 2))
    if scales is not None:
        assert scales.shape == (B, N) and oris.shape == (B, N)
        _f32c(scales, "scales"), _f32c(oris, "oris")
    if isinstance(cond, torch.Tensor):
        assert cond.shape == (B,)
        cond = cond.to(device=kpts.device, dtype=torch.float32).contiguous()
        _check(lib().lg_posenc_per_image(_ptr(kpts), _ptr(scales), _ptr(oris), _ptr(_f32c(size, "size")),
                                         _ptr(_f32c(wr, "wr")), _ptr(_f32c(wc, "wc")), _ptr(_f32c(bc, "bc")), _ptr(cond),
                                         _ptr(cos), _ptr(sin), B, N, _stream()), "lg_posenc_per_image")
    elif scales is not None:
        _check(lib().lg_posenc_scale_ori(_ptr(kpts), _ptr(scales), _ptr(oris), _ptr(_f32c(size, "size")),
                                         _ptr(_f32c(wr, "wr")), _ptr(_f32c(wc, "wc")), _ptr(_f32c(bc, "bc")), float(cond),
                                         _ptr(cos), _ptr(sin), B, N, _stream()), "lg_posenc_scale_ori")
    else:
        _check(lib().lg_posenc(_ptr(kpts), _ptr(_f32c(size, "size")), _ptr(_f32c(wr, "wr")), _ptr(_f32c(wc, "wc")),
                               _ptr(_f32c(bc, "bc")), float(cond), _ptr(cos), _ptr(sin), B, N, _stream()), "lg_posenc")
    _count(1)
    return cos, sin


def alloc_tiles(B, H, N, device):
    return torch.empty(B, H, tiles_pad(N), TILE_BYTES, device=device, dtype=torch.uint8)


def pack_self_qkv(qkv, cos, sin, lens=None, H=4):
    B, N, F = qkv.shape
    assert F == H * 192
    q, k, v = (alloc_tiles(B, H, N, qkv.device) for _ in range(3))
    _check(lib().lg_pack_self_qkv(_ptr(_f32c(qkv, "qkv")), _ptr(cos), _ptr(sin), _ptr(_i32(lens)), _ptr(q), _ptr(k),
                                  _ptr(v), B, N, H, tiles_pad(N), _stream()), "lg_pack_self_qkv")
    _count(1)
    return q, k, v


def pack_cross_qkv(qk, v, lens=None, H=4):
    B, N, F = qk.shape
    assert F == H * 64 and v.shape == qk.shape
    qt, vt = alloc_tiles(B, H, N, qk.device), alloc_tiles(B, H, N, qk.device)
    _check(lib().lg_pack_cross_qkv(_ptr(_f32c(qk, "qk")), _ptr(_f32c(v, "v")), _ptr(_i32(lens)), _ptr(qt), _ptr(vt),
                                   B, N, H, tiles_pad(N), _stream()), "lg_pack_cross_qkv")
    _count(1)
    return qt, vt


def attention(q_tiles, k_tiles, v_tiles, nq, nk, q_len=None, k_len=None, scale=0.125):
    """-> context (B, nq, H*64) f32."""
    B, H = q_tiles.shape[:2]
    out = torch.empty(B, nq, H * 64, device=q_tiles.device, dtype=torch.float32)
    if ATTN_EVENTS is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    _check(lib().lg_attention(_ptr(q_tiles), _ptr(k_tiles), _ptr(v_tiles), _ptr(out), _ptr(_i32(q_len)),
                              _ptr(_i32(k_len)), B, H, nq, nk, q_tiles.shape[2], k_tiles.shape[2], float(scale),
                              _stream()), "lg_attention")
    if ATTN_EVENTS is not None:
        e1.record()
        ATTN_EVENTS.append((e0, e1, 4.0 * 64 * nq * nk * B * H))  # algorithmic: QK^T + PV, 2 FLOP/MAC
    _count(1)
    return out


def assign(md0, md1, z0, z1, m_len=None, n_len=None, prescale=0.25):
    """md0 (B,M,256), md1 (B,N,256), z0 (B,M), z1 (B,N) f32 -> log assignment (B,M+1,N+1) f32."""
    B, M, D = md0.shape
    N = md1.shape[1]
    assert D == 256 and md1.shape[2] == 256
    scores = torch.empty(B, M + 1, N + 1, device=md0.device, dtype=torch.float32)
    ws = torch.empty(lib().lg_assign_ws_bytes(B, M, N), device=md0.device, dtype=torch.uint8)
    _check(lib().lg_assign(_ptr(_f32c(md0, "md0")), _ptr(_f32c(md1, "md1")), _ptr(_f32c(z0, "z0")),
                           _ptr(_f32c(z1, "z1")), _ptr(_i32(m_len)), _ptr(_i32(n_len)), float(prescale),
                           _ptr(scores), _ptr(ws), B, M, N, _stream()), "lg_assign")
    _count(5)
    return scores


def filter_matches(scores, th: float, m_len=None, n_len=None):
    """scores (B, M+1, N+1) f32, contiguous or a view of pitched rows (alloc_scores) -> m0 (B,M) i64, m1 (B,N) i64,
    ms0, ms1 f32."""
    assert scores.is_cuda and scores.dtype == torch.float32 and scores.stride(2) == 1
    B, M1, N1 = scores.shape
    assert scores.stride(0) == M1 * scores.stride(1), "image pitch = (M+1) * row pitch"
    M, N = M1 - 1, N1 - 1
    dev = scores.device
    m0 = torch.empty(B, M, device=dev, dtype=torch.int64)
    m1 = torch.empty(B, N, device=dev, dtype=torch.int64)
    ms0 = torch.empty(B, M, device=dev, dtype=torch.float32)
    ms1 = torch.empty(B, N, device=dev, dtype=torch.float32)
    if M == 0 or N == 0:
        raise RuntimeError("filter_matches: empty inner matrix (the reference's max() raises too)")
    ws = torch.empty(lib().lg_filter_ws_bytes(B, M, N), device=dev, dtype=torch.uint8)
    _check(lib().lg_filter_matches(_ptr(scores), scores.stride(1), _ptr(_i32(m_len)), _ptr(_i32(n_len)), float(th), _ptr(m0),
                                   _ptr(m1), _ptr(ms0), _ptr(ms1), _ptr(ws), B, M, N, _stream()), "lg_filter_matches")
    _count(2)
    return m0, m1, ms0, ms1


FILTER_STATIC_MAX_KEYS = 131072  # LG_FILTER_STATIC_MAX_KEYS of include/lg_b200.h


def filter_matches_static(scores, th: float, out, m_len=None, n_len=None):
    """lg_filter_matches_static: the caller passes the outputs `out` = (m0, m1, ms0, ms1); the call itself allocates
    nothing (key scratch = a static device array of the library; B*(M+N) <= FILTER_STATIC_MAX_KEYS)."""
    _f32c(scores, "scores")
    B, M1, N1 = scores.shape
    M, N = M1 - 1, N1 - 1
    m0, m1, ms0, ms1 = out
    assert m0.shape == (B, M) and m1.shape == (B, N) and m0.dtype == torch.int64 and m1.dtype == torch.int64
    assert ms0.shape == (B, M) and ms1.shape == (B, N) and ms0.dtype == torch.float32 and ms1.dtype == torch.float32
    assert all(t.is_cuda and t.is_contiguous() for t in out)
    _check(lib().lg_filter_matches_static(_ptr(scores), N1, _ptr(_i32(m_len)), _ptr(_i32(n_len)), float(th), _ptr(m0),
                                          _ptr(m1), _ptr(ms0), _ptr(ms1), B, M, N, _stream()),
           "lg_filter_matches_static")
    _count(2)
    return out


def assign_and_filter(md0, md1, z0, z1, th: float, m_len=None, n_len=None, prescale=0.25):
    """lg_assign_keys + lg_filter_from_keys: log assignment (B,M+1,N+1) and the filter_matches outputs, with the row /
    column argmax gathered inside the assignment kernel (no second pass over the matrix)."""
    B, M, D = md0.shape
    N = md1.shape[1]
    assert D == 256 and md1.shape[2] == 256
    dev = md0.device
    scores = torch.empty(B, M + 1, N + 1, device=dev, dtype=torch.float32)
    ws = torch.empty(lib().lg_assign_ws_bytes(B, M, N), device=dev, dtype=torch.uint8)
    keys = torch.empty(lib().lg_filter_ws_bytes(B, M, N), device=dev, dtype=torch.uint8)
    ml, nl = _i32(m_len), _i32(n_len)
    _check(lib().lg_assign_keys(_ptr(_f32c(md0, "md0")), _ptr(_f32c(md1, "md1")), _ptr(_f32c(z0, "z0")),
                                _ptr(_f32c(z1, "z1")), _ptr(ml), _ptr(nl), float(prescale), _ptr(scores), _ptr(ws),
                                _ptr(keys), B, M, N, _stream()), "lg_assign_keys")
    m0 = torch.empty(B, M, device=dev, dtype=torch.int64)
    m1 = torch.empty(B, N, device=dev, dtype=torch.int64)
    ms0 = torch.empty(B, M, device=dev, dtype=torch.float32)
    ms1 = torch.empty(B, N, device=dev, dtype=torch.float32)
    _check(lib().lg_filter_from_keys(_ptr(keys), _ptr(ml), _ptr(nl), float(th), _ptr(m0), _ptr(m1), _ptr(ms0),
                                     _ptr(ms1), B, M, N, _stream()), "lg_filter_from_keys")
    _count(6)
    return scores, (m0, m1, ms0, ms1)


def logit_rows(x, w, bias):
    """x (..., 256) f32 -> <x, w> + bias over the last axis (the matchability logits, lightglue.py:313-314)."""
    assert x.shape[-1] == 256
    out = torch.empty(x.shape[:-1], device=x.device, dtype=torch.float32)
    _check(lib().lg_logit_rows(_ptr(_f32c(x, "x")), _ptr(_f32c(w, "w")), _ptr(bias), _ptr(out), out.numel(), _stream()),
           "lg_logit_rows")
    _count(1)
    return out


def scores_pitch(n: int) -> int:
    """Row pitch (floats) of a log_assignment buffer with n+1 columns: a multiple of 32, so that every row and every
    64-byte segment the assignment kernel stores starts on a 128-byte line."""
    return (n + 1 + 31) // 32 * 32


def alloc_scores(B: int, M: int, N: int, device):
    """-> (storage (B, M+1, pitch), the (B, M+1, N+1) view callers see).  The reference's contract is shape and dtype
    (lightglue.py:567-577); the padding columns are never written."""
    buf = torch.empty(B, M + 1, scores_pitch(N), device=device, dtype=torch.float32)
    return buf, buf[:, :, : N + 1]


def assign_tiles(a_tiles, b_tiles, tpa, tpb, z0, z1, th=None, m_len=None, n_len=None, scale=1.0 / 16):
    """lg_assign_tiles: fp16 tiles of final_proj(desc0 / desc1) (linear_f16; image b at row tile b*tpa / b*tpb), z0 (B,M),
    z1 (B,N) -> log assignment (B,M+1,N+1) (a view of rows pitched to 128 bytes); with `th` also the filter_matches
    outputs (keys gathered while the values are written)."""
    B, M = z0.shape
    N = z1.shape[1]
    dev = z0.device
    buf, scores = alloc_scores(B, M, N, dev)
    ws = torch.empty(lib().lg_assign_tiles_ws_bytes(B, M, N), device=dev, dtype=torch.uint8)
    keys = None if th is None else torch.empty(lib().lg_filter_ws_bytes(B, M, N), device=dev, dtype=torch.uint8)
    ml, nl = _i32(m_len), _i32(n_len)
    _check(lib().lg_assign_tiles(_ptr(a_tiles), _ptr(b_tiles), int(tpa), int(tpb), _ptr(_f32c(z0, "z0")),
                                 _ptr(_f32c(z1, "z1")), _ptr(ml), _ptr(nl), float(scale), _ptr(buf), buf.shape[2],
                                 _ptr(ws), _ptr(keys), B, M, N, _stream()), "lg_assign_tiles")
    _count(4)
    if th is None:
        return scores
    m0 = torch.empty(B, M, device=dev, dtype=torch.int64)
    m1 = torch.empty(B, N, device=dev, dtype=torch.int64)
    ms0 = torch.empty(B, M, device=dev, dtype=torch.float32)
    ms1 = torch.empty(B, N, device=dev, dtype=torch.float32)
    _check(lib().lg_filter_from_keys(_ptr(keys), _ptr(ml), _ptr(nl), float(th), _ptr(m0), _ptr(m1), _ptr(ms0),
                                     _ptr(ms1), B, M, N, _stream()), "lg_filter_from_keys")
    _count(1)
    return scores, (m0, m1, ms0, ms1)


def token_scores(desc, w_tok, b_tok, w_match, b_match, lens=None):
    """-> (conf or None, match), each (B,N) f32. w_tok None skips the confidence head."""
    B, N, D = desc.shape
    assert D == 256
    match = torch.empty(B, N, device=desc.device, dtype=torch.float32)
    conf = torch.empty_like(match) if w_tok is not None else None
    _check(lib().lg_token_scores(_ptr(_f32c(desc, "desc")), _ptr(w_tok), _ptr(b_tok), _ptr(_f32c(w_match, "w_match")),
                                 _ptr(b_match), _ptr(_i32(lens)), _ptr(conf), _ptr(match), B, N, _stream()),
           "lg_token_scores")
    _count(1)
    return conf, match


def prune_plan(conf, match, thr: float, min_match: float, lens=None):
    """-> keep_idx (B,N) i32 (first counts[b,0] entries valid, ascending), counts (B,2) i32."""
    B, N = match.shape
    keep_idx = torch.empty(B, N, device=match.device, dtype=torch.int32)
    counts = torch.empty(B, 2, device=match.device, dtype=torch.int32)
    _check(lib().lg_prune_plan(_ptr(conf), _ptr(_f32c(match, "match")), _ptr(_i32(lens)), float(thr), float(min_match),
                               _ptr(keep_idx), _ptr(counts), B, N, _stream()), "lg_prune_plan")
    _count(1)
    return keep_idx, counts


def prune_gather(desc, cos, sin, ind, keep_idx, counts, prune, zero=False):
    """Gather kept rows into fresh (B,N,..) buffers (rows >= counts[b,0] are unspecified, or zero with zero=True:
    batches whose pairs keep different counts feed those rows to the layer kernels as padding)."""
    B, N, _ = desc.shape
    new = torch.zeros_like if zero else torch.empty_like
    desc_o, cos_o, sin_o, ind_o = new(desc), new(cos), new(sin), new(ind)
    assert ind.dtype == torch.int64 and prune.dtype == torch.int64
    _check(lib().lg_prune_gather(_ptr(_f32c(desc, "desc")), _ptr(_f32c(cos, "cos")), _ptr(_f32c(sin, "sin")), _ptr(ind),
                                 _ptr(keep_idx), _ptr(counts), _ptr(desc_o), _ptr(cos_o), _ptr(sin_o), _ptr(ind_o),
                                 _ptr(prune), B, N, N, prune.shape[1], _stream()), "lg_prune_gather")
    _count(1)
    return desc_o, cos_o, sin_o, ind_o


def score_prune_tiles(x, rot, x_out, rot_out, ind, ind_out, prune, prune_out, lens, lens_out, counts, w_tok, b_tok, w_match,
                      b_match, thr: float, min_match: float, do_prune: bool, B: int, n0: int, n1: int, tp0: int, tp1: int,
                      n0_orig: int, n1_orig: int):
    """lg_score_prune_tiles: ind / ind_out / prune / prune_out / lens / lens_out are (side 0, side 1) pairs (entries may
    be None where the header allows NULL); counts (B,4) i32."""
    pp = lambda pair, i: None if pair is None or pair[i] is None else pair[i].data_ptr()
    ws = torch.empty(lib().lg_score_prune_ws_bytes(B, tp0, tp1) // 8, device=x.device, dtype=torch.int64)
    _check(lib().lg_score_prune_tiles(_ptr(x), _ptr(rot), _ptr(x_out), _ptr(rot_out), pp(ind, 0), pp(ind, 1), pp(ind_out, 0),
                                      pp(ind_out, 1), pp(prune, 0), pp(prune, 1), pp(prune_out, 0), pp(prune_out, 1),
                                      pp(lens, 0), pp(lens, 1), pp(lens_out, 0), pp(lens_out, 1), _ptr(counts), _ptr(w_tok),
                                      _ptr(b_tok), _ptr(w_match), _ptr(b_match), float(thr), float(min_match),
                                      int(bool(do_prune)), _ptr(ws), B, n0, n1, tp0, tp1, n0_orig, n1_orig, _stream()),
           "lg_score_prune_tiles")
    _count(1)


# ----------------------------------------------------------------------------- split tiles + linears
SPLIT_ATOM = 32768


def alloc_bf16_tiles(row_tiles: int, k: int, device):
    """Plain 16-bit activation tiles [row_tiles][k/64][16 KB] (bf16 hi-only GEMM operand, or the fp16 tiles of linear_f16)."""
    return torch.empty(row_tiles, k // 64, TILE_BYTES, device=device, dtype=torch.uint8)


def alloc_split(row_tiles: int, k: int, device):
    """Split-tile activation buffer [row_tiles][k/64][hi 16 KB | lo 16 KB]."""
    return torch.empty(row_tiles, k // 64, SPLIT_ATOM, device=device, dtype=torch.uint8)


def pack_weight(w, nt: int = 256):
    """nn.Linear weight (n_out, k) f32 -> packed split-bf16 tiles (one-off, at weight-load time)."""
    n_out, k = w.shape
    out = torch.empty(n_out * k * 4, device=w.device, dtype=torch.uint8)
    _check(lib().lg_pack_weight(_ptr(_f32c(w, "w")), _ptr(out), n_out, k, nt, _stream()), "lg_pack_weight")
    _count(1)
    return out


def rows_to_tiles(x, tp: int, out=None):
    """x (B,N,K) f32 -> split tiles (B*tp, K/64, 32 KB)."""
    B, N, K = x.shape
    if out is None:
        out = alloc_split(B * tp, K, x.device)
    _check(lib().lg_rows_to_tiles(_ptr(_f32c(x, "x")), _ptr(out), B, N, K, tp, _stream()), "lg_rows_to_tiles")
    _count(1)
    return out


def tiles_to_rows(tiles, B: int, N: int, K: int, tp: int):
    x = torch.empty(B, N, K, device=tiles.device, dtype=torch.float32)
    _check(lib().lg_tiles_to_rows(_ptr(tiles), _ptr(x), B, N, K, tp, _stream()), "lg_tiles_to_rows")
    _count(1)
    return x


def rotary_tiles(cos, sin, tp: int, out=None):
    B, N, _ = cos.shape
    if out is None:
        out = torch.empty(B * tp, SPLIT_ATOM, device=cos.device, dtype=torch.uint8)
    _check(lib().lg_rotary_tiles(_ptr(_f32c(cos, "cos")), _ptr(_f32c(sin, "sin")), _ptr(out), B, N, tp, _stream()),
           "lg_rotary_tiles")
    _count(1)
    return out


def ln_gelu_tiles(h, gamma, beta, eps: float, out=None):
    rows = h.shape[0]
    assert h.shape[1] == 512
    if out is None:
        out = alloc_split(rows // TILE, 512, h.device)
    _check(lib().lg_ln_gelu_tiles(_ptr(_f32c(h, "h")), _ptr(_f32c(gamma, "gamma")), _ptr(_f32c(beta, "beta")), float(eps),
                                  _ptr(out), rows, _stream()), "lg_ln_gelu_tiles")
    _count(1)
    return out


def ffn_ln_gelu(a0, w, bias, gamma, beta, eps: float, a1=None, out=None):
    """GELU(LayerNorm([a0 | a1] . w^T + bias)) -> split tiles (row_tiles, 8, 32 KB); one fused launch."""
    rt = a0.shape[0]
    if out is None:
        out = alloc_split(rt, 512, a0.device)
    _check(lib().lg_ffn_ln_gelu(_ptr(a0), a0.shape[1], _ptr(a1), 0 if a1 is None else a1.shape[1], _ptr(w),
                                _ptr(_f32c(bias, "bias")), _ptr(_f32c(gamma, "gamma")), _ptr(_f32c(beta, "beta")),
                                float(eps), _ptr(out), rt, _stream()), "lg_ffn_ln_gelu")
    _count(1)
    return out


def linear_split(a0, w, bias, n_out: int, a1=None, resid=None, out=None, a1_hi_only=False):
    """[a0 | a1] . w^T + bias (+ resid) -> split tiles (row_tiles, n_out/64, 32 KB)."""
    rt = a0.shape[0]
    if out is None:
        out = alloc_split(rt, n_out, a0.device)
    _check(lib().lg_linear_split(_ptr(a0), a0.shape[1], _ptr(a1), 0 if a1 is None else a1.shape[1], int(a1_hi_only), _ptr(w),
                                 _ptr(_f32c(bias, "bias")), _ptr(resid), _ptr(out), rt, n_out, _stream()),
           "lg_linear_split")
    _count(1)
    return out


def linear_f16(a0, w, bias, n_out: int, a1=None, out=None, a1_hi_only=False):
    """[a0 | a1] . w^T + bias -> plain fp16 tiles (row_tiles, n_out/64, 16 KB): the operands of assign_tiles."""
    rt = a0.shape[0]
    if out is None:
        out = alloc_bf16_tiles(rt, n_out, a0.device)
    _check(lib().lg_linear_f16(_ptr(a0), a0.shape[1], _ptr(a1), 0 if a1 is None else a1.shape[1], int(a1_hi_only), _ptr(w),
                               _ptr(_f32c(bias, "bias")), _ptr(out), rt, n_out, _stream()), "lg_linear_f16")
    _count(1)
    return out


def linear_f32(a0, w, bias, n_out: int, a1=None, out=None, a1_hi_only=False):
    """[a0 | a1] . w^T + bias -> f32 rows (row_tiles*128, n_out)."""
    rt = a0.shape[0]
    if out is None:
        out = torch.empty(rt * TILE, n_out, device=a0.device, dtype=torch.float32)
    _check(lib().lg_linear_f32(_ptr(a0), a0.shape[1], _ptr(a1), 0 if a1 is None else a1.shape[1], int(a1_hi_only), _ptr(w),
                               _ptr(_f32c(bias, "bias")), _ptr(out), out.stride(0), rt, n_out, _stream()),
           "lg_linear_f32")
    _count(1)
    return out


def linear_qkv_self(a, w, bias, rot, B: int, tp0: int, tp1: int, bufs, H: int = 4):
    """bufs = ((q0,k0,v0), (q1,k1,v1) or None): attention operand tiles per side."""
    s0, s1 = bufs[0], (bufs[1] if tp1 else (None, None, None))
    _check(lib().lg_linear_qkv_self(_ptr(a), _ptr(w), _ptr(_f32c(bias, "bias")), _ptr(rot), _ptr(s0[0]), _ptr(s0[1]),
                                    _ptr(s0[2]), _ptr(s1[0]), _ptr(s1[1]), _ptr(s1[2]), B, H, tp0, tp1, _stream()),
           "lg_linear_qkv_self")
    _count(1)


def linear_qkv_cross(a, w, bias, B: int, tp0: int, tp1: int, bufs, H: int = 4):
    """bufs = ((qk0,v0), (qk1,v1) or None)."""
    s0, s1 = bufs[0], (bufs[1] if tp1 else (None, None))
    _check(lib().lg_linear_qkv_cross(_ptr(a), _ptr(w), _ptr(_f32c(bias, "bias")), _ptr(s0[0]), _ptr(s0[1]), _ptr(s1[0]),
                                     _ptr(s1[1]), B, H, tp0, tp1, _stream()), "lg_linear_qkv_cross")
    _count(1)


def attention_tiles(q_tiles, k_tiles, v_tiles, out_tiles, rt_base: int, nq, nk, q_len=None, k_len=None, scale=0.125,
                    cross=False):
    """Context written as split tiles (K atom = head) into out_tiles starting at row tile rt_base.
    cross: one of the two directions of CrossBlock; its ALGORITHMIC flops count the shared similarity once
    (6*d*M*N per layer for both directions), although this launch issues its own QK^T."""
    B, H = q_tiles.shape[:2]
    if ATTN_EVENTS is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    _check(lib().lg_attention_tiles(_ptr(q_tiles), _ptr(k_tiles), _ptr(v_tiles), _ptr(out_tiles), rt_base,
                                    _ptr(_i32(q_len)), _ptr(_i32(k_len)), B, H, nq, nk, q_tiles.shape[2],
                                    k_tiles.shape[2], float(scale), _stream()), "lg_attention_tiles")
    if ATTN_EVENTS is not None:
        e1.record()
        ATTN_EVENTS.append((e0, e1, (3.0 if cross else 4.0) * 64 * nq * nk * B * H))
    _count(1)


def attention_tiles2(side0, side1, out_tiles, scale=0.125, cross=False):
    """Both attention problems of a block in one launch; side = (q_tiles, k_tiles, v_tiles, rt_base, nq, nk, q_len,
    k_len).  cross: the two directions of a CrossBlock (algorithmic flops count the shared similarity once)."""
    B, H = side0[0].shape[:2]
    if ATTN_EVENTS is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    args = []
    for q, k, v, rt, nq, nk, ql, kl in (side0, side1):
        args += [_ptr(q), _ptr(k), _ptr(v), int(rt), _ptr(_i32(ql)), _ptr(_i32(kl)), int(nq), int(nk), q.shape[2], k.shape[2]]
    _check(lib().lg_attention_tiles2(*args, _ptr(out_tiles), B, H, float(scale), _stream()), "lg_attention_tiles2")
    if ATTN_EVENTS is not None:
        e1.record()
        flops = sum((3.0 if cross else 4.0) * 64 * sd[4] * sd[5] * B * H for sd in (side0, side1))
        ATTN_EVENTS.append((e0, e1, flops))
    _count(1)


# ----------------------------------------------------------------------------- fp32 path (lg_fp32.cu)
def gemm_f32(A, B, C, M, N, K, lda, ldb, ldc, *, bias=None, resid=None, ldr=0, b_kn=False, alpha=1.0, batch=1, inner=1,
             strides=(0,) * 8, m_len=None, k_len=None, offs=(0, 0, 0, 0)):
    """C = alpha * A . op(B) + bias + resid on raw strided views: A, B, C, resid are the BASE tensors (any shape, f32
    cuda), `offs` = element offsets of (A, B, C, resid) inside them, `strides` = (outer, inner) element strides of the
    batch index for A, B, C, resid.  See lg_gemm_f32 in include/lg_b200.h."""
    st = (_c.c_int64 * 8)(*[int(v) for v in strides])
    pa, pb, pc = A.data_ptr() + 4 * offs[0], B.data_ptr() + 4 * offs[1], C.data_ptr() + 4 * offs[2]
    pr = None if resid is None else resid.data_ptr() + 4 * offs[3]
    _check(lib().lg_gemm_f32(pa, pb, pc, _ptr(bias), pr, M, N, K, lda, ldb, ldc, ldr, int(b_kn), float(alpha), batch,
                             inner, st, _ptr(_i32(m_len)), _ptr(_i32(k_len)), _stream()), "lg_gemm_f32")
    _count(1)


def softmax_rows_f32(s, m_len=None, n_len=None, n=None):
    """s (B,H,M,pitch) f32, in place over the first n (<= pitch) columns."""
    B, H, M, pitch = s.shape
    _check(lib().lg_softmax_rows_f32(_ptr(_f32c(s, "s")), _ptr(_i32(m_len)), _ptr(_i32(n_len)), B, H, M,
                                     pitch if n is None else n, pitch, _stream()), "lg_softmax_rows_f32")
    _count(1)


def rotary_rows_f32(x, cos, sin, q_off, k_off, heads):
    rows, ld = x.shape[0] * x.shape[1], x.shape[2]
    _check(lib().lg_rotary_rows_f32(_ptr(_f32c(x, "x")), _ptr(_f32c(cos, "cos")), _ptr(_f32c(sin, "sin")), rows, ld,
                                    q_off, k_off, heads, _stream()), "lg_rotary_rows_f32")
    _count(1)


def ln_gelu_rows_f32(x, gamma, beta, eps):
    width = x.shape[-1]
    _check(lib().lg_ln_gelu_rows_f32(_ptr(_f32c(x, "x")), _ptr(_f32c(gamma, "gamma")), _ptr(_f32c(beta, "beta")),
                                     float(eps), x.numel() // width, width, _stream()), "lg_ln_gelu_rows_f32")
    _count(1)


# ----------------------------------------------------------------------------- nearest-neighbour matcher (lg_nn.cu)
def nn_assign(d0, d1, m_len=None, n_len=None):
    """d0 (B,M,256), d1 (B,N,256) -> similarity (B,M,N), log_assignment (B,M+1,N+1) (zero dustbins)."""
    B, M, D = d0.shape
    N = d1.shape[1]
    assert D == 256 and d1.shape[2] == 256, "the similarity kernel is built for 256-d descriptors"
    alloc = torch.zeros if m_len is not None or n_len is not None else torch.empty
    sim = alloc(B, M, N, device=d0.device, dtype=torch.float32)
    scores = torch.empty(B, M + 1, N + 1, device=d0.device, dtype=torch.float32)
    ws = torch.empty(lib().lg_assign_ws_bytes(B, M, N), device=d0.device, dtype=torch.uint8)
    _check(lib().lg_nn_assign(_ptr(_f32c(d0, "d0")), _ptr(_f32c(d1, "d1")), _ptr(_i32(m_len)), _ptr(_i32(n_len)),
                              _ptr(sim), _ptr(scores), _ptr(ws), B, M, N, _stream()), "lg_nn_assign")
    _count(5)
    return sim, scores


def nn_match(sim, ratio_thresh=None, distance_thresh=None, mutual=True, m_len=None, n_len=None):
    """find_nn both ways + mutual_check on sim (B,M,N) -> m0 (B,M) i64, m1 (B,N) i64, ms0, ms1 f32."""
    B, M, N = sim.shape
    dev = sim.device
    m0 = torch.empty(B, M, device=dev, dtype=torch.int64)
    m1 = torch.empty(B, N, device=dev, dtype=torch.int64)
    ms0 = torch.empty(B, M, device=dev, dtype=torch.float32)
    ms1 = torch.empty(B, N, device=dev, dtype=torch.float32)
    ws = torch.empty(lib().lg_nn_ws_bytes(B, M, N), device=dev, dtype=torch.uint8)
    _check(lib().lg_nn_match(_ptr(_f32c(sim, "sim")), _ptr(_i32(m_len)), _ptr(_i32(n_len)), float(ratio_thresh or 0.0),
                             float(distance_thresh or 0.0), int(bool(mutual)), _ptr(m0), _ptr(m1), _ptr(ms0), _ptr(ms1),
                             _ptr(ws), B, M, N, _stream()), "lg_nn_match")
    _count(3)
    return m0, m1, ms0, ms1
