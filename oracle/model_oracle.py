"""CPU oracle for the ProteinBERT pretraining forward / loss / backward.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this file; it
is used by ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` as the checker and the timed CPU
baseline.  It is a channels-last, functional restatement (plain torch CPU ops,
any float dtype) of the reference algorithm; every function cites the reference
lines it follows (paths relative to the upstream repo root).

Parity is pinned: ``tests/golden/make_golden.py`` imports the real reference
``ProteinBERT/modules.py`` (it needs only torch) in the build container, runs
it on seeded inputs and stores inputs / weights / outputs / loss / gradients in
``tests/golden/model_tiny.npz``; ``tests/test_oracle_golden.py`` checks this
restatement against those vectors.

Parameters are passed as a flat ``dict[str, Tensor]`` keyed exactly like the
reference ``state_dict`` (133 tensors at the default config), plus the
attention-head tensors, which the reference keeps in a plain python list and
therefore never registers (``modules.py:73-81``); here they are keyed
``proteinBERT_blocks.{i}.global_attention_layer.global_attention_heads.{h}.W{q,k,v}_parameter``.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

HEAD_FMT = "proteinBERT_blocks.{i}.global_attention_layer.global_attention_heads.{h}.{n}_parameter"


def gelu(t: Tensor) -> Tensor:
    """Exact-erf GELU (``nn.GELU()`` default, used everywhere in modules.py)."""
    return 0.5 * t * (1.0 + torch.erf(t * (1.0 / math.sqrt(2.0))))


def conv_same(x: Tensor, w: Tensor, b: Tensor, dilation: int) -> Tensor:
    """'same' cross-correlation on channels-last ``x:[B,L,Ci]`` with the
    reference weight layout ``w:[Co,Ci,k]`` (``nn.Conv1d`` at modules.py:124-147).

    out[b,l,co] = b[co] + sum_t sum_ci w[co,ci,t] * x[b, l+(t-k//2)*d, ci]
    (zero outside the sequence, no kernel flip).
    """
    B, L, Ci = x.shape
    Co, _, k = w.shape
    half = (k // 2) * dilation
    xp = F.pad(x, (0, 0, half, half))
    out = b.view(1, 1, Co).expand(B, L, Co).clone()
    for t in range(k):
        out = out + xp[:, t * dilation:t * dilation + L, :] @ w[:, :, t].t()
    return out


def ln_joint(u: Tensor, gamma: Tensor, beta: Tensor, eps: float = 1e-5) -> Tensor:
    """``nn.LayerNorm((L, C))``: one mean / biased variance per sample over the
    L*C elements, affine ``[L,C]`` (modules.py:148-151,161-164,212,217)."""
    mean = u.mean(dim=(1, 2), keepdim=True)
    var = ((u - mean) ** 2).mean(dim=(1, 2), keepdim=True)
    return (u - mean) * torch.rsqrt(var + eps) * gamma + beta


def ln_rows(u: Tensor, gamma: Tensor, beta: Tensor, eps: float = 1e-5) -> Tensor:
    """``nn.LayerNorm(G)`` on the global track (modules.py:183-186,196-199)."""
    mean = u.mean(dim=-1, keepdim=True)
    var = ((u - mean) ** 2).mean(dim=-1, keepdim=True)
    return (u - mean) * torch.rsqrt(var + eps) * gamma + beta


def attention_as_executed(h: Tensor, g: Tensor, heads, w_param: Tensor) -> Tensor:
    """GlobalAttention exactly as the reference executes it
    (modules.py:49-60 per head, :87-92 for the head mix).

    The global vector is repeated ``key_dim`` times, so Q has K identical rows;
    ``Softmax(dim=1)`` then normalises over that repeated axis (not over the
    sequence), which makes every probability exactly 1/K.
    """
    outs = []
    for (wq, wk, wv) in heads:
        K = wq.shape[1]
        xg = g.unsqueeze(1).expand(-1, K, -1)                 # [B,K,G]  (:51)
        q = torch.tanh(xg @ wq)                               # [B,K,K]  (:53)
        kt = torch.tanh(h @ wk)                               # [B,L,K]  (:54)
        v = gelu(h @ wv)                                      # [B,L,Dv] (:55)
        s = (q @ kt.transpose(1, 2)) / math.sqrt(float(K))    # [B,K,L]  (:58)
        p = torch.softmax(s, dim=1)                           # over the K axis (:34)
        outs.append(p @ v)                                    # [B,K,Dv]
    cat = torch.cat(outs, dim=2)                              # [B,K,G]  (:92)
    return torch.matmul(w_param, cat)                         # [B,G]


def attention_closed_form(h: Tensor, g: Tensor, heads, w_param: Tensor) -> Tensor:
    """Same function evaluated in closed form: (sum(W_parameter)/K) * sum_l gelu(h_l Wv_all).
    An independent second oracle for the attention layer."""
    K = heads[0][0].shape[1]
    wv_all = torch.cat([hd[2] for hd in heads], dim=1)        # [C,G]
    pool = gelu(h @ wv_all).sum(dim=1)                        # [B,G]
    return pool * (w_param.sum() / K)


def attention_sequence(h: Tensor, g: Tensor, heads, w_param: Tensor) -> Tensor:
    """The reference with its softmax moved to the sequence axis (``dim=2``): the
    semantics the paper describes.  Not the default anywhere; kept so that a
    ``softmax_axis='sequence'`` mode has something to be checked against."""
    outs = []
    for (wq, wk, wv) in heads:
        K = wq.shape[1]
        xg = g.unsqueeze(1).expand(-1, K, -1)
        q = torch.tanh(xg @ wq)
        kt = torch.tanh(h @ wk)
        v = gelu(h @ wv)
        s = (q @ kt.transpose(1, 2)) / math.sqrt(float(K))
        p = torch.softmax(s, dim=2)
        outs.append(p @ v)
    return torch.matmul(w_param, torch.cat(outs, dim=2))


_ATTENTION = {
    "as_executed": attention_as_executed,
    "closed_form": attention_closed_form,
    "sequence": attention_sequence,
}


def num_blocks_of(params: Dict[str, Tensor]) -> int:
    n = 0
    while f"proteinBERT_blocks.{n}.local_norm_1.weight" in params:
        n += 1
    return n


def heads_of(params: Dict[str, Tensor], i: int):
    heads = []
    h = 0
    while HEAD_FMT.format(i=i, h=h, n="Wq") in params:
        heads.append(tuple(params[HEAD_FMT.format(i=i, h=h, n=n)] for n in ("Wq", "Wk", "Wv")))
        h += 1
    return heads


def block_forward(params, i: int, x: Tensor, g: Tensor, dilation: int = 5,
                  attention: str = "as_executed") -> Tuple[Tensor, Tensor]:
    """One ProteinBERTBlock on channels-last ``x:[B,L,C]``, ``g:[B,G]``
    (modules.py:201-231; sums are taken left to right as written at :211)."""
    p = f"proteinBERT_blocks.{i}."
    nar = gelu(conv_same(x, params[p + "local_narrow_conv_layer.0.weight"],
                         params[p + "local_narrow_conv_layer.0.bias"], 1))           # :205
    wid = gelu(conv_same(x, params[p + "local_wide_conv_layer.0.weight"],
                         params[p + "local_wide_conv_layer.0.bias"], dilation))      # :206
    g2l = gelu(g @ params[p + "global_to_local_linear_layer.0.weight"].t()
               + params[p + "global_to_local_linear_layer.0.bias"])                  # :208
    u = x + nar + wid + g2l.unsqueeze(1)                                             # :209-211
    h = ln_joint(u, params[p + "local_norm_1.weight"], params[p + "local_norm_1.bias"])   # :212
    lin = gelu(h @ params[p + "local_linear_layer.0.weight"].t()
               + params[p + "local_linear_layer.0.bias"])                            # :214
    h = ln_joint(h + lin, params[p + "local_norm_2.weight"], params[p + "local_norm_2.bias"])  # :216-217
    att = _ATTENTION[attention](h, g, heads_of(params, i),
                                params[p + "global_attention_layer.W_parameter"])    # :219
    gl = gelu(g @ params[p + "global_linear_layer_1.0.weight"].t()
              + params[p + "global_linear_layer_1.0.bias"])                          # :221
    g1 = ln_rows(g + gl + att, params[p + "global_norm_1.weight"], params[p + "global_norm_1.bias"])  # :223-224
    gl2 = gelu(g1 @ params[p + "global_linear_layer_2.0.weight"].t()
               + params[p + "global_linear_layer_2.0.bias"])                         # :226
    g2 = ln_rows(g1 + gl2, params[p + "global_norm_2.weight"], params[p + "global_norm_2.bias"])      # :228-229
    return h, g2


def forward(params: Dict[str, Tensor], x_local: Tensor, x_global: Tensor,
            dilation: int = 5, attention: str = "as_executed") -> Tuple[Tensor, Tensor]:
    """ProteinBERT.forward (modules.py:295-304).  Returns the token probabilities
    ``[B,L,V]`` -- softmax over the BATCH axis, because ``nn.Softmax()`` without a
    dim on a 3-D input resolves to dim 0 (modules.py:283) -- and the annotation
    probabilities ``[B,A]`` (sigmoid, :292)."""
    x = params["local_embedding.weight"][x_local]                                    # :300
    g = gelu(x_global @ params["global_linear_layer.0.weight"].t()
             + params["global_linear_layer.0.bias"])                                 # :301
    for i in range(num_blocks_of(params)):
        x, g = block_forward(params, i, x, g, dilation, attention)
    logits = x @ params["pretraining_local_output.0.weight"].t() + params["pretraining_local_output.0.bias"]
    p_local = torch.softmax(logits, dim=0)                                           # :283 (implicit dim -> 0)
    p_global = torch.sigmoid(g @ params["pretraining_global_output.0.weight"].t()
                             + params["pretraining_global_output.0.bias"])           # :286-293
    return p_local, p_global


def pretraining_loss(p_local: Tensor, p_global: Tensor, y_local: Tensor, y_global: Tensor,
                     w_local: Tensor, w_global: Tensor) -> Tensor:
    """The loss expression of utils.py:293-294 with the loss modules of
    dummy_tests.py:132-133: CrossEntropyLoss(reduction='none') applied to the
    PROBABILITIES (a second, vocabulary-axis log-softmax) times float64 token
    weights, plus BCELoss(reduction='none') (logs clamped at -100) times float64
    annotation weights; both means divide by the element count.  float64 result."""
    lsm = p_local - torch.logsumexp(p_local, dim=2, keepdim=True)
    ce = -lsm.gather(2, y_local.unsqueeze(2)).squeeze(2)                             # [B,L]
    logp = torch.clamp(torch.log(p_global), min=-100.0)
    log1mp = torch.clamp(torch.log(1.0 - p_global), min=-100.0)
    bce = -(y_global * logp + (1.0 - y_global) * log1mp)
    return (ce * w_local).mean() + (bce * w_global).mean()


def loss_and_grads(params: Dict[str, Tensor], batch, dilation: int = 5, attention: str = "as_executed"):
    """fwd + loss + bwd on the CPU.  ``batch`` = (X_local i64[B,L], X_global f[B,A],
    Y_local i64[B,L], Y_global f[B,A], W_local f64[B,L], W_global f64[B,A]).
    Returns (p_local, p_global, loss, grads dict incl. head tensors)."""
    xl, xg, yl, yg, wl, wg = batch
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    p_local, p_global = forward(leaves, xl, xg, dilation, attention)
    loss = pretraining_loss(p_local, p_global, yl, yg, wl, wg)
    loss.backward()
    grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in leaves.items()}
    return p_local.detach(), p_global.detach(), loss.detach(), grads


def init_params(sequences_length: int, num_annotations: int, local_dim: int, global_dim: int,
                key_dim: int, num_heads: int, num_blocks: int, conv_kernel_size: int = 9,
                vocab_size: int = 26, seed: int = 0, dtype=torch.float32) -> Dict[str, Tensor]:
    """Random parameters with the reference's shapes and initial distributions
    (PyTorch defaults for Linear / Conv1d / LayerNorm / Embedding; ``torch.randn``
    for the head tensors and W_parameter, modules.py:36-47,82-85).  The draw ORDER
    is not the reference's; goldens carry the reference's own weights."""
    gen = torch.Generator().manual_seed(seed)
    L, A, C, G, K, H, V = sequences_length, num_annotations, local_dim, global_dim, key_dim, num_heads, vocab_size

    def uni(shape, fan_in):
        bound = 1.0 / math.sqrt(fan_in)
        return (torch.rand(shape, generator=gen, dtype=torch.float64) * 2 - 1).mul(bound).to(dtype)

    def randn(shape):
        return torch.randn(shape, generator=gen, dtype=torch.float64).to(dtype)

    p: Dict[str, Tensor] = {}
    p["local_embedding.weight"] = randn((V, C))
    p["global_linear_layer.0.weight"] = uni((G, A), A)
    p["global_linear_layer.0.bias"] = uni((G,), A)
    for i in range(num_blocks):
        b = f"proteinBERT_blocks.{i}."
        p[b + "global_attention_layer.W_parameter"] = randn((K,))
        for name, fan in (("local_narrow_conv_layer", C * conv_kernel_size), ("local_wide_conv_layer", C * conv_kernel_size)):
            p[b + name + ".0.weight"] = uni((C, C, conv_kernel_size), fan)
            p[b + name + ".0.bias"] = uni((C,), fan)
        p[b + "local_norm_1.weight"] = torch.ones(L, C, dtype=dtype)
        p[b + "local_norm_1.bias"] = torch.zeros(L, C, dtype=dtype)
        p[b + "local_linear_layer.0.weight"] = uni((C, C), C)
        p[b + "local_linear_layer.0.bias"] = uni((C,), C)
        p[b + "local_norm_2.weight"] = torch.ones(L, C, dtype=dtype)
        p[b + "local_norm_2.bias"] = torch.zeros(L, C, dtype=dtype)
        p[b + "global_to_local_linear_layer.0.weight"] = uni((C, G), G)
        p[b + "global_to_local_linear_layer.0.bias"] = uni((C,), G)
        for j in (1, 2):
            p[b + f"global_linear_layer_{j}.0.weight"] = uni((G, G), G)
            p[b + f"global_linear_layer_{j}.0.bias"] = uni((G,), G)
            p[b + f"global_norm_{j}.weight"] = torch.ones(G, dtype=dtype)
            p[b + f"global_norm_{j}.bias"] = torch.zeros(G, dtype=dtype)
        for h in range(H):
            p[HEAD_FMT.format(i=i, h=h, n="Wv")] = randn((C, G // H))
            p[HEAD_FMT.format(i=i, h=h, n="Wk")] = randn((C, K))
            p[HEAD_FMT.format(i=i, h=h, n="Wq")] = randn((G, K))
    p["pretraining_local_output.0.weight"] = uni((V, C), C)
    p["pretraining_local_output.0.bias"] = uni((V,), C)
    p["pretraining_global_output.0.weight"] = uni((A, G), G)
    p["pretraining_global_output.0.bias"] = uni((A,), G)
    return p


def synthetic_batch(B: int, L: int, A: int, seed: int = 0, vocab_size: int = 26):
    """Device-ready synthetic batch of the shapes/dtypes the reference dataset
    yields (data_processing.py:159-180) without going through the string path:
    random lengths (pad + full-length rows), 5 %-corrupted X tokens, Bernoulli(0.005)
    annotations, half of the samples with all annotations hidden."""
    gen = torch.Generator().manual_seed(seed)
    lengths = torch.randint(2, 2 * L, (B,), generator=gen).clamp(max=L)
    pos = torch.arange(L).unsqueeze(0)
    valid = pos < lengths.unsqueeze(1)
    y_local = torch.randint(4, vocab_size, (B, L), generator=gen)
    y_local[:, 0] = 1
    y_local[torch.arange(B), lengths - 1] = 2
    y_local = torch.where(valid, y_local, torch.zeros_like(y_local))
    corrupt = (torch.rand(B, L, generator=gen) < 0.05) & (y_local > 2)
    x_local = torch.where(corrupt, torch.randint(3, vocab_size, (B, L), generator=gen), y_local)
    y_global = (torch.rand(B, A, generator=gen) < 0.005).float()
    keep = (torch.rand(B, A, generator=gen) < 0.75).float()
    add = (torch.rand(B, A, generator=gen) < 1e-4).float()
    shown = (torch.rand(B, 1, generator=gen) > 0.5).float()
    x_global = (y_global + add) * keep * shown
    w_local = (y_local != 0).double()
    w_global = y_global.bool().any(dim=1, keepdim=True).double().expand(B, A).contiguous()
    return x_local, x_global, y_local, y_global, w_local, w_global
