"""Drop-in replacements for the four nn.Modules of the reference ``ProteinBERT/modules.py``.

Same class names, constructor signatures, dict-in/dict-out forwards, attribute names and ``state_dict``
keys (133 tensors at the default configuration); sub-modules are created in the reference's order, so the
same ``torch.manual_seed`` yields the same initial weights.  The arithmetic is done by the sm_100a kernels
of libpb_b200.so (see ops.py); the ``nn.Conv1d`` / ``nn.Linear`` / ``nn.LayerNorm`` objects below are only
parameter containers -- their own ``forward`` is never called.

Reference behaviours that are reproduced on purpose:
  * the attention heads live in a plain python list, so their tensors are NOT registered (modules.py:73-81);
  * ``Softmax(dim=1)`` in the head normalises over the repeated-query axis, which makes the layer
    ``(sum(W_parameter)/key_dim) * sum_l gelu(x_l Wv)`` (modules.py:34,51-60,92);
  * the token head's ``nn.Softmax()`` acts on the batch axis (modules.py:283).

Extra, optional constructor keywords (all default to the reference's behaviour):
  ``precision``     "fp32" (default; fp32 storage and fp32 arithmetic, matches the reference to ~1e-5) or "bf16" (bf16
                    activations, tcgen05 tensor-core contractions with fp32 accumulation, fp32/fp64 statistics);
  ``softmax_axis``  "reference" (default): the attention layer in the closed form the reference's softmax axis implies,
                    on the tensor cores; "reference_as_executed": the same function through the op-faithful fused kernel
                    (tanh(g Wq), tanh(h Wk), scores, softmax over the repeated-query axis, gelu(h Wv), P.V) -- a check of
                    the closed form; "sequence": that kernel with the softmax over the sequence positions, the paper's
                    semantics -- the only mode in which Wq and Wk influence the output;
  ``register_heads``  False (default, as the reference: heads in a plain list, their tensors unregistered) or True (heads in
                    an nn.ModuleList: they appear in parameters() / state_dict() and are trained).
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn

from . import ops
from ._lib import ACT_GELU, ACT_SIGMOID, BF16, F32

_DEFAULT_DEVICE = ('cuda' if torch.cuda.is_available() else 'cpu')
_DEFAULT_PRECISION = os.environ.get("PB_B200_PRECISION", "fp32")


def _axis(softmax_axis: str) -> int:
    """kernel code of the softmax axis: 0 = the reference's repeated-query axis, 1 = sequence positions"""
    if softmax_axis in ("reference", "reference_as_executed"):
        return 0
    if softmax_axis == "sequence":
        return 1
    raise ValueError(f"softmax_axis must be 'reference', 'reference_as_executed' or 'sequence', got {softmax_axis!r}")


def _dt(precision: str) -> int:
    if precision == "fp32":
        return F32
    if precision == "bf16":
        return BF16
    raise ValueError(f"precision must be 'fp32' or 'bf16', got {precision!r}")


class GlobalAttentionHead(nn.Module):
    """reference: modules.py:21-60."""

    def __init__(self, local_dim: int, global_dim: int, value_dim: int, key_dim: int,
                 device: torch.device = _DEFAULT_DEVICE, softmax_axis: str = "reference"):
        super().__init__()
        _axis(softmax_axis)
        self.softmax_axis = softmax_axis
        self.key_dim = key_dim
        self.global_dim = global_dim
        self.gelu = nn.GELU()
        self.tanh = nn.Tanh()
        self.softmax = nn.Softmax(dim=1)
        self.Wv_parameter = nn.Parameter(data=torch.randn(local_dim, value_dim, device=device), requires_grad=True)
        self.Wk_parameter = nn.Parameter(data=torch.randn(local_dim, key_dim, device=device), requires_grad=True)
        self.Wq_parameter = nn.Parameter(data=torch.randn(global_dim, key_dim, device=device), requires_grad=True)

    def forward(self, x):
        """{"local": [B,L,C], "global": [B,G]} -> [B, key_dim, value_dim].  Every one of the key_dim rows is
        (1/key_dim) * sum_l gelu(x_l Wv): the K query rows are identical copies of the global vector, and the
        softmax runs over that axis (modules.py:51-60), so Wq / Wk cancel out of the result exactly."""
        if self.softmax_axis == "reference":
            pool = ops.AttnPoolFn.apply(x["local"], self.Wv_parameter.t())          # [B, Dv]
            return (pool / self.key_dim).unsqueeze(1).expand(-1, self.key_dim, -1)
        # op-faithful kernel, one head; a unit mix (sum = 1) leaves the head's own [K, Dv] output, all rows equal
        mix = torch.full((self.key_dim,), 1.0 / self.key_dim, device=self.Wv_parameter.device)
        out = ops.AttentionFn.apply(x["local"], x["global"], self.Wq_parameter.unsqueeze(0), self.Wk_parameter.unsqueeze(0),
                                    self.Wv_parameter.unsqueeze(0), mix, _axis(self.softmax_axis))
        return out.unsqueeze(1).expand(-1, self.key_dim, -1)


class GlobalAttention(nn.Module):
    """reference: modules.py:63-92."""

    def __init__(self, num_heads: int, local_dim: int, global_dim: int, value_dim: int, key_dim: int,
                 device: torch.device = _DEFAULT_DEVICE, softmax_axis: str = "reference", register_heads: bool = False):
        super().__init__()
        _axis(softmax_axis)
        self.softmax_axis = softmax_axis
        # a plain list, exactly like the reference: the head tensors are not registered parameters (register_heads=True
        # is the opt-in fix: an nn.ModuleList)
        heads = [GlobalAttentionHead(local_dim=local_dim, global_dim=global_dim, value_dim=value_dim, key_dim=key_dim,
                                     device=device, softmax_axis=softmax_axis) for _ in range(num_heads)]
        self.global_attention_heads = nn.ModuleList(heads) if register_heads else heads
        self.W_parameter = nn.Parameter(data=torch.randn(key_dim, device=device), requires_grad=True)

    def wv_all_t(self) -> torch.Tensor:
        """cat_heads(Wv)^T, [G, C] (autograd routes the gradient back to each head's Wv_parameter)."""
        return torch.cat([h.Wv_parameter for h in self.global_attention_heads], dim=1).t()

    def stacked(self):
        """(Wq [H,G,K], Wk [H,C,K], Wv [H,C,Dv]): the heads' matrices stacked for the fused kernel (gradients flow back to each head)."""
        hs = self.global_attention_heads
        return (torch.stack([h.Wq_parameter for h in hs]), torch.stack([h.Wk_parameter for h in hs]),
                torch.stack([h.Wv_parameter for h in hs]))

    def forward(self, x):
        if self.softmax_axis == "reference":
            pool = ops.AttnPoolFn.apply(x["local"], self.wv_all_t())
            return pool * (self.W_parameter.sum() / self.W_parameter.numel())
        return ops.AttentionFn.apply(x["local"], x["global"], *self.stacked(), self.W_parameter, _axis(self.softmax_axis))


class ProteinBERTBlock(nn.Module):
    """reference: modules.py:95-231."""

    def __init__(self, sequences_length, local_dim: int, global_dim: int, num_heads: int, key_dim: int,
                 conv_kernel_size: int = 9, wide_conv_dilation: int = 5, device: torch.device = _DEFAULT_DEVICE,
                 precision: str = _DEFAULT_PRECISION, softmax_axis: str = "reference", register_heads: bool = False):
        super().__init__()
        _axis(softmax_axis)
        self.softmax_axis = softmax_axis
        assert global_dim % num_heads == 0, \
            f"Global_dim must be divisible by num_heads. Global_dim: {global_dim}. Num_heads: {num_heads}"
        self.sequences_length = sequences_length
        self.local_dim = local_dim
        self.conv_kernel_size = conv_kernel_size
        self.wide_conv_dilation = wide_conv_dilation
        self.precision = precision
        self.head_grads = True          # the reference's autograd also fills the (unregistered) heads' .grad
        self.fuse_global = os.environ.get("PB_B200_FUSE_GLOBAL", "1") == "1"   # 4-launch global track (ops_global_fused.cu)

        self.global_attention_layer = GlobalAttention(num_heads=num_heads, local_dim=local_dim, global_dim=global_dim,
                                                      value_dim=int(global_dim / num_heads), key_dim=key_dim, device=device,
                                                      softmax_axis=softmax_axis, register_heads=register_heads)
        self.local_narrow_conv_layer = nn.Sequential(
            nn.Conv1d(local_dim, local_dim, kernel_size=conv_kernel_size, stride=1, dilation=1, padding="same", device=device),
            nn.GELU())
        self.local_wide_conv_layer = nn.Sequential(
            nn.Conv1d(local_dim, local_dim, kernel_size=conv_kernel_size, stride=1, dilation=wide_conv_dilation,
                      padding="same", device=device),
            nn.GELU())
        self.local_norm_1 = nn.LayerNorm(normalized_shape=(sequences_length, local_dim), device=device)
        self.local_linear_layer = nn.Sequential(nn.Linear(local_dim, local_dim, device=device), nn.GELU())
        self.local_norm_2 = nn.LayerNorm(normalized_shape=(sequences_length, local_dim), device=device)
        self.global_to_local_linear_layer = nn.Sequential(nn.Linear(global_dim, local_dim, device=device), nn.GELU())
        self.global_linear_layer_1 = nn.Sequential(nn.Linear(global_dim, global_dim, device=device), nn.GELU())
        self.global_norm_1 = nn.LayerNorm(normalized_shape=global_dim, device=device)
        self.global_linear_layer_2 = nn.Sequential(nn.Linear(global_dim, global_dim, device=device), nn.GELU())
        self.global_norm_2 = nn.LayerNorm(normalized_shape=global_dim, device=device)

    def __getstate__(self):
        """torch.save(model) (utils.py:341) pickles the module: the per-step staging caches are rebuilt on demand, not saved."""
        st = dict(self.__dict__)
        for k in ("_staged", "_stage_table", "_prestaged"):
            st.pop(k, None)
        return st

    def local_track(self, xl: torch.Tensor, g2l: torch.Tensor, arena=None):
        """modules.py:205-219 without the global side: xl [B,L,C], g2l [B,C] -> (h [B,L,C], pool [B,G] = sum_l gelu(h Wv_all))."""
        if xl.shape[1] != self.sequences_length:
            raise ValueError(f"sequence length {xl.shape[1]} != sequences_length {self.sequences_length} "
                             "(the (L,C) LayerNorm fixes it, modules.py:148-151)")
        nconv, wconv = self.local_narrow_conv_layer[0], self.local_wide_conv_layer[0]
        lin = self.local_linear_layer[0]
        att = self.global_attention_layer
        dt = ops.dt_of(xl)
        if getattr(self, "_prestaged", None) == dt:          # ProteinBERT.forward staged all blocks in one launch
            staged, self._prestaged = self._staged, None
        else:
            staged = ops.stage_blocks(self, [self], dt)[0]
        return ops.LocalTrackFn.apply(                                                         # :205-219
            xl, g2l, nconv.weight, nconv.bias, wconv.weight, wconv.bias,
            self.local_norm_1.weight, self.local_norm_1.bias, lin.weight, lin.bias,
            self.local_norm_2.weight, self.local_norm_2.bias, staged,
            self.conv_kernel_size, self.wide_conv_dilation, self.head_grads, arena,
            *[hd.Wv_parameter for hd in att.global_attention_heads])

    def global_post_params(self):
        """(W_parameter, LN1 weight/bias, W4, b4, LN2 weight/bias): the post half of this block's global track (:92,:223-229)."""
        gl2 = self.global_linear_layer_2[0]
        return (self.global_attention_layer.W_parameter, self.global_norm_1.weight, self.global_norm_1.bias, gl2.weight, gl2.bias,
                self.global_norm_2.weight, self.global_norm_2.bias)

    def global_pre_params(self):
        """(W1, b1, W3, b3): global-to-local linear (:208) and global linear 1 (:221), both read the block's input g."""
        a, b = self.global_to_local_linear_layer[0], self.global_linear_layer_1[0]
        return (a.weight, a.bias, b.weight, b.bias)

    def forward_channels_last(self, xl: torch.Tensor, g: torch.Tensor, arena=None):
        """xl [B,L,C] in the activation dtype, g [B,G] float -> (h [B,L,C], g_out [B,G])."""
        if xl.shape[1] != self.sequences_length:
            raise ValueError(f"sequence length {xl.shape[1]} != sequences_length {self.sequences_length} "
                             "(the (L,C) LayerNorm fixes it, modules.py:148-151)")
        nconv, wconv = self.local_narrow_conv_layer[0], self.local_wide_conv_layer[0]
        lin = self.local_linear_layer[0]
        att = self.global_attention_layer
        g2l_lin = self.global_to_local_linear_layer[0]
        dt = ops.dt_of(xl)
        if getattr(self, "_prestaged", None) == dt:          # ProteinBERT.forward staged all blocks in one launch
            staged, self._prestaged = self._staged, None
        else:
            staged = ops.stage_blocks(self, [self], dt)[0]
        gl1, gl2 = self.global_linear_layer_1[0], self.global_linear_layer_2[0]
        fused = self.fuse_global and ops.global_fused_ok(g.shape[0], g.shape[1], g2l_lin.weight.shape[0])
        if fused:       # both linears that read the block's input g in one launch
            link = ops.GlobalLink()
            g2l, gh = ops.GlobalPreFn.apply(g, g2l_lin.weight, g2l_lin.bias, gl1.weight, gl1.bias, link)   # :208, :221
        else:
            g2l = ops.linear_act(g, g2l_lin.weight, g2l_lin.bias, ACT_GELU)                    # modules.py:208
        closed = self.softmax_axis == "reference"
        h, pool = ops.LocalTrackFn.apply(                                                      # :205-219
            xl, g2l, nconv.weight, nconv.bias, wconv.weight, wconv.bias,
            self.local_norm_1.weight, self.local_norm_1.bias, lin.weight, lin.bias,
            self.local_norm_2.weight, self.local_norm_2.bias, staged,
            self.conv_kernel_size, self.wide_conv_dilation, self.head_grads, arena,
            *([hd.Wv_parameter for hd in att.global_attention_heads] if closed else []))
        w_mix = att.W_parameter
        if not closed:
            # the attention layer through the op-faithful fused kernel (softmax over the reference's axis or over the
            # sequence): it returns sum(W_parameter) * P.V, so the kernels behind it get a unit mix (sum(w)/K = 1)
            pool = att({"local": h, "global": g})                                              # :219
            w_mix = torch.ones_like(att.W_parameter)
        if fused:       # LN(g + gh + att), Linear+GELU, LN in one launch                       # :223-229
            g2 = ops.GlobalPostFn.apply(g, gh, pool, w_mix, self.global_norm_1.weight, self.global_norm_1.bias,
                                        gl2.weight, gl2.bias, self.global_norm_2.weight, self.global_norm_2.bias, link)
            return h, g2
        gh = ops.linear_act(g, gl1.weight, gl1.bias, ACT_GELU)                                  # :221
        g1 = ops.ResLnRowsFn.apply(g, gh, pool, w_mix, self.global_norm_1.weight, self.global_norm_1.bias)   # :223-224
        gh2 = ops.linear_act(g1, gl2.weight, gl2.bias, ACT_GELU)                                # :226
        g2 = ops.ResLnRowsFn.apply(g1, gh2, None, None, self.global_norm_2.weight, self.global_norm_2.bias)           # :228-229
        return h, g2

    def forward(self, x):
        adt = ops.act_dtype(_dt(self.precision))
        xl = x["local"].permute(0, 2, 1)              # reference block I/O is channels-first [B,C,L]
        if xl.dtype != adt:
            xl = xl.to(adt)
        h, g = self.forward_channels_last(xl, x["global"].float())
        return {"local": h.permute(0, 2, 1), "global": g}


class ProteinBERT(nn.Module):
    """reference: modules.py:234-304."""

    def __init__(self, sequences_length: int, num_annotations: int, local_dim: int, global_dim: int, key_dim: int,
                 num_heads: int, num_blocks: int, conv_kernel_size: int = 9, wide_conv_dilation: int = 5,
                 vocab_size: int = 26, device: torch.device = _DEFAULT_DEVICE, precision: str = _DEFAULT_PRECISION,
                 softmax_axis: str = "reference", register_heads: bool = False):
        super().__init__()
        _dt(precision)
        _axis(softmax_axis)
        self.precision = precision
        self.softmax_axis = softmax_axis
        self.local_embedding = nn.Embedding(num_embeddings=vocab_size, embedding_dim=local_dim, device=device)
        self.global_linear_layer = nn.Sequential(nn.Linear(num_annotations, global_dim, device=device), nn.GELU())
        self.proteinBERT_blocks = nn.Sequential(*[
            ProteinBERTBlock(sequences_length=sequences_length, global_dim=global_dim, local_dim=local_dim,
                             key_dim=key_dim, num_heads=num_heads, conv_kernel_size=conv_kernel_size,
                             wide_conv_dilation=wide_conv_dilation, device=device, precision=precision,
                             softmax_axis=softmax_axis, register_heads=register_heads)
            for _ in range(num_blocks)
        ])
        self.pretraining_local_output = nn.Sequential(nn.Linear(local_dim, vocab_size, device=device), nn.Softmax())
        self.pretraining_global_output = nn.Sequential(nn.Linear(global_dim, num_annotations, device=device), nn.Sigmoid())

    def __getstate__(self):
        st = dict(self.__dict__)
        for k in ("_stage_table", "_boundary", "_boundary_at", "_wo_pads"):
            st.pop(k, None)
        return st

    def set_precision(self, precision: str):
        _dt(precision)
        self.precision = precision
        for blk in self.proteinBERT_blocks:
            blk.precision = precision
        return self

    def set_head_grads(self, flag: bool):
        """Whether backward also produces .grad for the unregistered attention-head tensors (the reference
        does; no optimizer ever reads them).  True by default."""
        for blk in self.proteinBERT_blocks:
            blk.head_grads = bool(flag)
        return self

    def uses_cluster_units(self, batch: int) -> bool:
        """Whether forward runs the global track as one cluster launch per block boundary (ops.GTrackFn) at this batch size."""
        blocks = list(self.proteinBERT_blocks)
        G = self.global_linear_layer[0].weight.shape[0]
        return (all(b.fuse_global and b.softmax_axis == "reference" for b in blocks) and os.environ.get("PB_B200_GTRACK", "1") == "1"
                and ops.gtrack_ok(batch, G, blocks[0].local_dim))

    def head_tensors(self):
        """The unregistered Wq/Wk/Wv tensors, keyed like oracle/model_oracle.py."""
        out = {}
        for i, blk in enumerate(self.proteinBERT_blocks):
            for h, head in enumerate(blk.global_attention_layer.global_attention_heads):
                for n in ("Wq", "Wk", "Wv"):
                    out[f"proteinBERT_blocks.{i}.global_attention_layer.global_attention_heads.{h}.{n}_parameter"] = \
                        getattr(head, f"{n}_parameter")
        return out

    def forward(self, x):
        """{"local": int64 [B,L], "global": float [B,A]} -> (token probs [B,L,V], annotation probs [B,A])."""
        dt = _dt(self.precision)
        x_local, x_global = x["local"], x["global"]
        ops._need_cuda(x_local, x_global, self.local_embedding.weight)      # no CPU path: refuse before touching any stream
        blocks = list(self.proteinBERT_blocks)
        lo, go = self.pretraining_local_output[0], self.pretraining_global_output[0]
        # every block's conv packs / W2 / Wv -- and the token head's padded output matrix -- in ONE launch.  (Issued on the side
        # stream beside the embedding and the annotation-input layer it measured 7 us SLOWER per step: fork + join cost more
        # than the 21 us kernel hides.)
        extra, wo_pads = (), None
        Bsz0, Lseq0 = x_local.shape
        if dt == ops.BF16 and bool(ops.lib().pb_token_head_tc_ok(Bsz0, Lseq0, lo.weight.shape[1], lo.weight.shape[0])) and lo.weight.shape[1] == 128:
            pads = getattr(self, "_wo_pads", None)
            if pads is None or pads.device != lo.weight.device:
                pads = torch.zeros((2, 128, 128), dtype=torch.bfloat16, device=lo.weight.device)     # rows >= V stay zero
                self._wo_pads = pads
            V_ = lo.weight.shape[0]
            extra = ((lo.weight, pads[0].data_ptr(), pads[1].data_ptr(), 0, V_, 128, 0, 128, 128),)
            wo_pads = (pads[0], pads[1])
        ops.stage_blocks(self, blocks, dt, extra)
        h = ops.EmbeddingFn.apply(x_local, self.local_embedding.weight, dt)                     # modules.py:300
        gin = self.global_linear_layer[0]
        tc = dt == ops.BF16           # annotation layers: tensor cores in bf16 precision, fp32 SIMT otherwise
        g = ops.linear_act(x_global.float(), gin.weight, gin.bias, ACT_GELU, tc)                # :301
        # training.GraphedPretrainStep: where backward is split in two (an int: the boundary tensors are kept in _boundary) or
        # marked for the gradient exchange (a collection of block indices: ops.CUT_HOOK fires at each in backward)
        cut = getattr(self, "_boundary_at", None)
        cuts = () if cut is None else ((cut,) if isinstance(cut, int) else tuple(cut))
        Bsz, Lseq, Cloc = h.shape
        arena = ops.ZeroArena(len(blocks) * ops.ZeroArena.block_bytes(Bsz, Lseq, Cloc, g.shape[1]), h.device)   # one fill per step
        Cdim = blocks[0].local_dim
        use_units = self.uses_cluster_units(Bsz)
        if use_units:
            # the global track runs as one cluster launch per block boundary: [post of block i ; pre of block i+1]
            links = [ops.GTrackLink() for _ in range(len(blocks) + 1)]
            none7 = (None,) * 7
            _, g2l, gh = ops.GTrackFn.apply(g, None, None, *none7, *blocks[0].global_pre_params(), Cdim, None, links[0])
            for i, blk in enumerate(blocks):
                blk._prestaged = dt
                h, pool = blk.local_track(h, g2l, arena)
                if i + 1 in cuts:
                    # a proper cut of the autograd graph: everything above consumes only these four tensors (the inputs of the
                    # unit that spans the boundary and of block `cut`'s local track); h and pool come out of the same node
                    h, pool, g, gh = ops.CutFn.apply(i + 1, h, pool, g, gh)
                    if isinstance(cut, int):
                        self._boundary = (h, pool, g, gh)
                nxt = blocks[i + 1].global_pre_params() if i + 1 < len(blocks) else (None,) * 4
                g, g2l, gh = ops.GTrackFn.apply(g, gh, pool, *blk.global_post_params(), *nxt, Cdim, links[i], links[i + 1])
        else:
            for i, blk in enumerate(blocks):
                if i in cuts:
                    h, g = ops.CutFn.apply(i, h, g)
                    if isinstance(cut, int):
                        self._boundary = (h, g)
                blk._prestaged = dt
                h, g = blk.forward_channels_last(h, g, arena)
        p_local = ops.TokenHeadFn.apply(h, lo.weight, lo.bias, wo_pads)                         # :277-284
        p_global = ops.linear_act(g, go.weight, go.bias, ACT_SIGMOID, tc)                       # :286-293
        return p_local, p_global
