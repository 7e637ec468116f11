"""Host-side driver of the transformer layers on split-tile activations.

Mirrors TransformerLayer.forward / SelfBlock.forward / CrossBlock.forward of the reference
(gluefactory/models/matchers/lightglue.py:178-191,220-249,258-281) as a fixed sequence of C-ABI
kernel launches; torch only owns the device buffers and the stream.

Both images of a pair share every weight of a layer (lightglue.py:268-272), so their tokens are
processed as ONE activation matrix: rows = [image-0 batch (B*tp0 tiles) | image-1 batch (B*tp1 tiles)].

Per block (self or cross):
    1. lg_linear_qkv_{self,cross}   x -> bf16 attention operand tiles (head split, rotary fused)
    2. lg_attention_tiles2          context of image 0 and image 1 (one launch) -> split tiles (K atom = head)
    3. lg_ffn_ln_gelu               GELU(LayerNorm([x | ctx] . W0'^T + b0')) -> split tiles (one fused launch)
    4. lg_linear_split              x + h . W3^T + b3 -> next x (split tiles)
Step 3 uses the out_proj / to_out projection folded into ffn.0 (exact algebra, done once in fp64):
    ffn0(cat[x, Wo ctx + bo]) = W0x x + (W0m Wo) ctx + (W0m bo + b0)          (lightglue.py:190-191,246-248)
"""
import torch

from . import _cabi as C


def _qkv_self_perm(heads: int, dh: int, device):
    """reference Wqkv output row h*3dh + d*3 + part (lightglue.py:185) -> packed row h*3dh + part*dh + d."""
    h = torch.arange(heads, device=device)[:, None, None]
    part = torch.arange(3, device=device)[None, :, None]
    d = torch.arange(dh, device=device)[None, None, :]
    return (h * 3 * dh + d * 3 + part).reshape(-1)


class PackedWeights:
    """Device-side packed (split-bf16, swizzled) copies of the module's weights; rebuilt when any
    parameter changes (load_state_dict, .to(device))."""

    def __init__(self, model):
        self.signature = self.sign(model)
        dev = next(model.parameters()).device
        d = model.conf.descriptor_dim
        perm = _qkv_self_perm(model.conf.num_heads, d // model.conf.num_heads, dev)
        self.layers = []
        for layer in model.transformers:
            sa, ca = layer.self_attn, layer.cross_attn
            blk_s = self._ffn(sa.ffn, sa.out_proj, d)
            blk_s["wqkv"] = C.pack_weight(sa.Wqkv.weight.detach()[perm].contiguous().float(), 192)
            blk_s["bqkv"] = sa.Wqkv.bias.detach()[perm].contiguous().float()
            blk_c = self._ffn(ca.ffn, ca.to_out, d)
            blk_c["wqkv"] = C.pack_weight(torch.cat([ca.to_qk.weight, ca.to_v.weight], 0).detach().contiguous().float())
            blk_c["bqkv"] = torch.cat([ca.to_qk.bias, ca.to_v.bias], 0).detach().contiguous().float()
            self.layers.append((blk_s, blk_c))
        self.final = [(C.pack_weight(h.final_proj.weight.detach().contiguous().float()),
                       h.final_proj.bias.detach().contiguous().float()) for h in model.log_assignment]
        self.input_proj = None
        if model.conf.input_dim != d:
            self.input_proj = (C.pack_weight(model.input_proj.weight.detach().contiguous().float()),
                               model.input_proj.bias.detach().contiguous().float())

    @staticmethod
    def sign(model):
        """(address, version) of every parameter: a load_state_dict, a .to(device) or an in-place update re-packs.  The
        Parameter objects of a module persist, so their list is cached (walking model.parameters() costs ~0.3 ms per
        forward: a fifth of a one-pair call)."""
        plist = model.__dict__.get("_lg_params")
        if plist is None:
            plist = model.__dict__["_lg_params"] = list(model.parameters())
        return tuple((p.data_ptr(), p._version) for p in plist)

    @staticmethod
    def _ffn(ffn, proj, d):
        w0, b0 = ffn[0].weight.detach().double(), ffn[0].bias.detach().double()
        wo, bo = proj.weight.detach().double(), proj.bias.detach().double()
        w0m = w0[:, d:]
        w0f = torch.cat([w0[:, :d], w0m @ wo], 1).float().contiguous()
        b0f = (b0 + w0m @ bo).float().contiguous()
        return {
            "w0": C.pack_weight(w0f), "b0": b0f,
            "ln_g": ffn[1].weight.detach().float().contiguous(), "ln_b": ffn[1].bias.detach().float().contiguous(),
            "eps": float(ffn[1].eps),
            "w3": C.pack_weight(ffn[3].weight.detach().contiguous().float()), "b3": ffn[3].bias.detach().float().contiguous(),
        }


class TileEngine:
    """Activation buffers for one (B, M, N) problem + the per-layer launch sequence."""

    def __init__(self, packed: PackedWeights, B: int, M: int, N: int, device, heads: int = 4):
        self.pk, self.B, self.M, self.N, self.H, self.dev = packed, B, M, N, heads, device
        self.tp0, self.tp1 = C.tiles_pad(M), C.tiles_pad(N)
        self.rt0 = B * self.tp0
        self.rt = B * (self.tp0 + self.tp1)
        self.x = C.alloc_split(self.rt, 256, device)
        self.x_alt = C.alloc_split(self.rt, 256, device)
        self.ctx = C.alloc_split(self.rt, 256, device)
        self.h = C.alloc_split(self.rt, 512, device)
        self.rot = torch.empty(self.rt, C.SPLIT_ATOM, device=device, dtype=torch.uint8)
        self.qkv = [tuple(C.alloc_tiles(B, heads, n, device) for _ in range(3)) for n in (M, N)]
        # CUDA-graph state of the layer stack (run_layers): static length buffers, the captured graph
        self.len0 = torch.empty(B, device=device, dtype=torch.int32)
        self.len1 = torch.empty(B, device=device, dtype=torch.int32)
        self.graphs = {}
        self.eager_runs = 0

    # ------------------------------------------------------------------ loading
    def load(self, desc0, desc1, cs0, cs1):
        """desc (B,n,D_in) f32 rows, cs = compact (cos, sin) tables (B,n,32)."""
        C.rotary_tiles(cs0[0], cs0[1], self.tp0, out=self.rot[: self.rt0])
        C.rotary_tiles(cs1[0], cs1[1], self.tp1, out=self.rot[self.rt0:])
        if self.pk.input_proj is None:
            self.load_rows(desc0, desc1)
        else:  # input_proj (lightglue.py:370-373,478-479)
            d_in = desc0.shape[-1]
            raw = C.alloc_split(self.rt, d_in, self.dev)
            C.rows_to_tiles(desc0, self.tp0, out=raw[: self.rt0])
            C.rows_to_tiles(desc1, self.tp1, out=raw[self.rt0:])
            w, b = self.pk.input_proj
            C.linear_split(raw, w, b, 256, out=self.x)

    def load_rows(self, desc0, desc1, cs0=None, cs1=None):
        """256-d descriptors (already projected) as rows; optionally new rotary tables (after pruning)."""
        if cs0 is not None:
            C.rotary_tiles(cs0[0], cs0[1], self.tp0, out=self.rot[: self.rt0])
            C.rotary_tiles(cs1[0], cs1[1], self.tp1, out=self.rot[self.rt0:])
        C.rows_to_tiles(desc0, self.tp0, out=self.x[: self.rt0])
        C.rows_to_tiles(desc1, self.tp1, out=self.x[self.rt0:])

    # ------------------------------------------------------------------ one transformer layer
    def _ffn(self, blk):
        C.ffn_ln_gelu(self.x, blk["w0"], blk["b0"], blk["ln_g"], blk["ln_b"], blk["eps"], a1=self.ctx, out=self.h)
        C.linear_split(self.h, blk["w3"], blk["b3"], 256, resid=self.x, out=self.x_alt)
        self.x, self.x_alt = self.x_alt, self.x

    def layer(self, i: int, len0=None, len1=None):
        blk_s, blk_c = self.pk.layers[i]
        B, M, N = self.B, self.M, self.N
        (q0, k0, v0), (q1, k1, v1) = self.qkv
        scale = 64 ** -0.5
        # SelfBlock on both images (lightglue.py:268-269)
        C.linear_qkv_self(self.x, blk_s["wqkv"], blk_s["bqkv"], self.rot, B, self.tp0, self.tp1, self.qkv, self.H)
        C.attention_tiles2((q0, k0, v0, 0, M, M, len0, len0), (q1, k1, v1, self.rt0, N, N, len1, len1), self.ctx, scale)
        self._ffn(blk_s)
        # CrossBlock (lightglue.py:270-272): shared to_qk projection, both directions
        C.linear_qkv_cross(self.x, blk_c["wqkv"], blk_c["bqkv"], B, self.tp0, self.tp1, ((q0, v0), (q1, v1)), self.H)
        C.attention_tiles2((q0, q1, v1, 0, M, N, len0, len1), (q1, q0, v0, self.rt0, N, M, len1, len0), self.ctx, scale,
                           cross=True)
        self._ffn(blk_c)

    def run_layers(self, n_layers: int, len0=None, len1=None, use_graph: bool = True):
        """All transformer layers.  The 10 x n_layers launches are captured once per engine into a CUDA graph
        (streams + graphs instead of a tracing compiler) and replayed: the layer stack only touches this
        engine's own buffers, so the graph is valid for every later batch of the same shape."""
        if len0 is not None:
            self.len0.copy_(len0)
            self.len1.copy_(len1)
        l0, l1 = (self.len0, self.len1) if len0 is not None else (None, None)
        if not use_graph or C.ATTN_EVENTS is not None:
            for i in range(n_layers):
                self.layer(i, l0, l1)
            return
        key = (n_layers, len0 is not None)
        if key in self.graphs and self.graphs[key] is None:  # capture failed once for this shape: plain launches
            for i in range(n_layers):
                self.layer(i, l0, l1)
            return
        g = self.graphs.get(key)
        if g is None:
            if self.eager_runs == 0:
                # first use: eager, so every kernel has set its attributes / loaded before any capture
                self.eager_runs = 1
                for i in range(n_layers):
                    self.layer(i, l0, l1)
                return
            x_before, n0 = self.x, C.LAUNCHES
            graph = torch.cuda.CUDAGraph()
            try:
                # thread_local: a CUDA call of ANOTHER host thread during the capture (a DataLoader pin_memory
                # thread of the reference's export / eval loops) must not abort it
                with C.pinned_stream(None), torch.cuda.graph(graph, capture_error_mode="thread_local"):
                    for i in range(n_layers):  # (the capture runs on its own stream: no pinned handle)
                        self.layer(i, l0, l1)
            except RuntimeError:
                # capture failed (e.g. an allocation raced it): this shape runs as plain launches from now on
                C.LAUNCHES = n0
                self.graphs[key] = None
                torch.cuda.synchronize()
                for i in range(n_layers):
                    self.layer(i, l0, l1)
                return
            assert self.x is x_before, "an even number of buffer swaps per layer keeps x in place"
            g = self.graphs[key] = (graph, C.LAUNCHES - n0)
            C.LAUNCHES = n0  # capture launched nothing; replays account for the kernels below
        g[0].replay()
        C._count(g[1])

    # ------------------------------------------------------------------ adaptive width (lg_score_prune_tiles)
    def rot_alt(self):
        """Second rotary-table buffer: the compaction of a layer writes x_alt / rot_alt, swap_pruned() makes them current."""
        if getattr(self, "_rot_alt", None) is None:
            self._rot_alt = torch.empty_like(self.rot)
        return self._rot_alt

    def swap_pruned(self):
        self.x, self.x_alt = self.x_alt, self.x
        self.rot, self._rot_alt = self._rot_alt, self.rot

    # ------------------------------------------------------------------ read-out
    def rows(self):
        """Current descriptors as f32 rows: (B,M,256), (B,N,256)."""
        return (C.tiles_to_rows(self.x[: self.rt0], self.B, self.M, 256, self.tp0),
                C.tiles_to_rows(self.x[self.rt0:], self.B, self.N, 256, self.tp1))

    def final_proj_tiles(self, i: int):
        """MatchAssignment.final_proj of head i (lightglue.py:310) as fp16 tiles, the operand format of lg_assign_tiles:
        -> (tiles of image side 0, tiles of side 1), image b at row tile b*tp0 / b*tp1."""
        w, b = self.pk.final[i]
        if getattr(self, "md", None) is None:
            self.md = C.alloc_bf16_tiles(self.rt, 256, self.dev)
        C.linear_f16(self.x, w, b, 256, out=self.md)
        return self.md[: self.rt0], self.md[self.rt0:]
