"""torch.autograd bindings of the C-ABI kernels.

PyTorch is plumbing here: it owns device memory (caching allocator), the current stream and the autograd
graph; every forward/backward below is one or more calls into libpb_b200.so with raw device pointers.
There is no eager fallback -- tensors that are not on a CUDA device are refused.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import ACT_GELU, ACT_NONE, ACT_SIGMOID, BF16, F32, check, lib

LN_EPS = 1e-5


def _p(t):
    return None if t is None else t.data_ptr()


def _st():
    return torch.cuda.current_stream().cuda_stream


PROFILE = None      # bench.py: {"pb_<op>": []} -> (start, end) CUDA event pairs around that op's launches


def _call(name, *args):
    """One C-ABI call: raises PBError on a non-zero status; optionally bracketed by CUDA events."""
    fn = getattr(lib(), name)
    if PROFILE is not None and name in PROFILE:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rc = fn(*args)
        b.record()
        PROFILE[name].append((a, b))
    else:
        rc = fn(*args)
    check(rc, name)


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.PBError("proteinbert_b200 ops run on a CUDA device (sm_100a) only; got a CPU tensor. "
                               "There is no CPU fallback.")


def act_dtype(dt: int):
    return torch.bfloat16 if dt == BF16 else torch.float32


def dt_of(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return BF16
    if t.dtype == torch.float32:
        return F32
    raise _lib.PBError(f"unsupported activation dtype {t.dtype}")


# ---- gradient sink -------------------------------------------------------------------------------------------------
# training.FlatGradBucket lays all registered gradients out in ONE flat buffer (what the data-parallel all-reduce sends).
# While a sink is active, the backward functions below write each parameter's gradient straight into its slice of that
# buffer (fresh views, so autograd adopts them as .grad without a copy) instead of a torch.empty tensor: no gather / scatter
# pass around the collective.  Slices documented as accumulated ("zero on entry") are cleared by the bucket once per step.
GRAD_SINK = None


class GradSink:
    def __init__(self, flat, table):
        self.flat, self.table = flat, table            # table: parameter data_ptr -> (offset, shape)

    def take(self, key, shape):
        ent = self.table.get(key)
        if ent is None or tuple(ent[1]) != tuple(shape):
            return None
        off, shp = ent
        n = 1
        for d in shp:
            n *= int(d)
        return self.flat[off:off + n].view(shp)


def _gout(key, shape, device, zero=False):
    """Gradient output for the parameter whose storage starts at `key` (data_ptr): its bucket slice when a sink is active."""
    if GRAD_SINK is not None and key is not None:
        v = GRAD_SINK.take(key, shape)
        if v is not None:
            return v
    return (torch.zeros if zero else torch.empty)(shape, dtype=torch.float32, device=device)


_WS = {}


def _workspace(nbytes: int, device) -> torch.Tensor:
    """Per-(device, stream) split-K scratch, grown on demand (kernels never allocate)."""
    key = (device, _st())
    buf = _WS.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _WS[key] = buf
    return buf


def _wgrad_ws(device):
    ws = _workspace(lib().pb_workspace_bytes(b"wgrad", 0, 0, 0), device)
    return ws.data_ptr(), ws.numel()


def to_wt(w: torch.Tensor, dt: int) -> torch.Tensor:
    """fp32 master weight -> the storage type the big contractions read (`wt` in pb_b200.h)."""
    w = w.contiguous()
    if dt == F32:
        return w
    out = torch.empty(w.shape, dtype=torch.bfloat16, device=w.device)
    _call("pb_cast_f32_to_bf16", _p(w), _p(out), w.numel(), _st())
    return out


def conv_pack(w: torch.Tensor, dt: int):
    """[Co,Ci,k] -> ([k,Co,Ci], [k,Ci,Co]) in the `wt` type."""
    Co, Ci, k = w.shape
    w = w.contiguous()
    fwd = torch.empty((k, Co, Ci), dtype=act_dtype(dt), device=w.device)
    bwd = torch.empty((k, Ci, Co), dtype=act_dtype(dt), device=w.device)
    _call("pb_conv_weight_pack", _p(w), _p(fwd), _p(bwd), Co, k, dt, _st())
    return fwd, bwd


class StagedBlockWeights:
    """Persistent staging buffers of one block's local-track weights in the `wt` storage type: the tap-major conv
    packs, W2 / W2^T and Wv_all / Wv_all^T.  Filled every forward by ONE pb_stage_weights launch for all blocks
    (see stage_blocks); the fp32 masters stay in the nn.Module."""

    def __init__(self, block, dt):
        nconv, wconv = block.local_narrow_conv_layer[0], block.local_wide_conv_layer[0]
        lin = block.local_linear_layer[0]
        heads = block.global_attention_layer.global_attention_heads
        C, _, k = nconv.weight.shape
        Dv = heads[0].Wv_parameter.shape[1]
        G = Dv * len(heads)
        dev, adt = nconv.weight.device, act_dtype(dt)
        e = lambda *shape: torch.empty(shape, dtype=adt, device=dev)
        self.dt = dt
        self.wn_f, self.wn_b, self.ww_f, self.ww_b = e(k, C, C), e(k, C, C), e(k, C, C), e(k, C, C)
        self.W2w, self.W2Tw = e(C, C), e(C, C)
        self.wv, self.wvT = e(C, G), e(G, C)
        self.srcs = [nconv.weight, wconv.weight, lin.weight] + [h.Wv_parameter for h in heads]
        es = self.wv.element_size()
        self.jobs = [(nconv.weight, self.wn_f.data_ptr(), self.wn_b.data_ptr(), 1, C, C, k, 0, 0),
                     (wconv.weight, self.ww_f.data_ptr(), self.ww_b.data_ptr(), 1, C, C, k, 0, 0),
                     (lin.weight, self.W2w.data_ptr(), self.W2Tw.data_ptr(), 0, C, C, 0, C, C)]
        for h, head in enumerate(heads):
            self.jobs.append((head.Wv_parameter, self.wv.data_ptr() + h * Dv * es, self.wvT.data_ptr() + h * Dv * C * es,
                              0, C, Dv, 0, G, C))

    def key(self):
        return (self.dt,) + tuple(t.data_ptr() for t in self.srcs)


_STAGE_DTYPE = [("src", "<u8"), ("dst0", "<u8"), ("dst1", "<u8"), ("kind", "<i4"), ("d0", "<i4"), ("d1", "<i4"),
                ("d2", "<i4"), ("ld0", "<i4"), ("ld1", "<i4"), ("reserved", "<i4"), ("pad", "<i4")]   # 56 B = sizeof(pb_stage_desc)


def stage_blocks(owner, blocks, dt, extra_jobs=()):
    """Stage the local-track weights of `blocks` with one kernel launch.  The device job table is cached on
    `owner` and rebuilt only when a parameter's storage or the precision changes.  extra_jobs: more rows of the same
    table, (src, dst0_ptr, dst1_ptr, kind, d0, d1, d2, ld0, ld1) -- the token head's padded output matrix."""
    import numpy as np
    staged = []
    for blk in blocks:
        sw = getattr(blk, "_staged", None)
        if sw is None or sw.key() != (dt,) + tuple(t.data_ptr() for t in sw.srcs):
            sw = StagedBlockWeights(blk, dt)
            blk._staged = sw
        staged.append(sw)
    key = tuple(sw.key() for sw in staged) + tuple((j[0].data_ptr(),) + tuple(j[1:]) for j in extra_jobs)
    cache = getattr(owner, "_stage_table", None)
    if cache is None or cache[0] != key:
        rows = []
        for sw in staged:
            for (src, d0p, d1p, kind, a, b, c, ld0, ld1) in sw.jobs:
                if not src.is_contiguous():
                    raise _lib.PBError("parameters must be contiguous")
                rows.append((src.data_ptr(), d0p, d1p, kind, a, b, c, ld0, ld1, 0, 0))
        for (src, d0p, d1p, kind, a, b, c, ld0, ld1) in extra_jobs:
            if not src.is_contiguous():
                raise _lib.PBError("parameters must be contiguous")
            rows.append((src.data_ptr(), d0p, d1p, kind, a, b, c, ld0, ld1, 0, 0))
        arr = np.array(rows, dtype=_STAGE_DTYPE)
        table = torch.from_numpy(arr.view(np.uint8).copy()).to(staged[0].wv.device)
        cache = (key, table, len(rows))
        owner._stage_table = cache
    _need_cuda(cache[1])
    _call("pb_stage_weights", _p(cache[1]), cache[2], dt, _st())
    return staged


# ------------------------------------------------------------------------------------------------
class EmbeddingFn(torch.autograd.Function):
    """nn.Embedding forward / embedding_dense_backward (modules.py:249-253,300)."""

    @staticmethod
    def forward(ctx, tok, table, dt):
        _need_cuda(tok, table)
        tok = tok.contiguous()
        B, L = tok.shape
        V, C = table.shape
        out = torch.empty((B, L, C), dtype=act_dtype(dt), device=table.device)
        _call("pb_embedding_fwd", _p(tok), _p(table.contiguous()), _p(out), B * L, C, V, dt, _st())
        ctx.save_for_backward(tok)
        ctx.shape = (V, C, dt)
        ctx.key = table.data_ptr()
        return out

    @staticmethod
    def backward(ctx, dout):
        (tok,) = ctx.saved_tensors
        V, C, dt = ctx.shape
        dout = dout.contiguous()
        dtable = _gout(ctx.key, (V, C), dout.device)
        # (on the side stream, beside the last global-track unit, this kernel -- 296 CTAs with 100 KB of shared memory each --
        # kept the unit's clusters off the SMs: 28 us slower per step)
        _call("pb_embedding_bwd", _p(tok), _p(dout), _p(dtable), tok.numel(), C, V, dt, _st())
        return None, dtable, None


class LinearActFn(torch.autograd.Function):
    """act(X W^T + b) on the global track, float32 tensors (modules.py:166-199,255-262,286-293).  tc=True takes the
    tcgen05 path (bf16 operands, fp32 accumulation) when the shape allows: the annotation layers in bf16 precision."""

    @staticmethod
    def forward(ctx, x, W, b, act, tc=False):
        _need_cuda(x, W, b)
        x = x.contiguous().float()
        W = W.contiguous()
        M, K = x.shape
        N = W.shape[0]
        y = torch.empty((M, N), dtype=torch.float32, device=x.device)
        z = torch.empty_like(y) if act == ACT_GELU else None
        tc = bool(tc) and bool(lib().pb_linear_tc_ok(M, N, K))
        if tc:
            ws = _workspace(lib().pb_linear_tc_workspace(M, N, K), x.device)
            _call("pb_linear_fwd_tc", _p(x), _p(W), _p(b), _p(z), _p(y), M, N, K, act, _p(ws), ws.numel(), _st())
        else:
            nbytes = lib().pb_workspace_bytes(b"linear_fwd", M, N, K)
            ws = _workspace(nbytes, x.device)
            _call("pb_linear_fwd", _p(x), _p(W), _p(b), _p(z), _p(y), M, N, K, act, _p(ws), ws.numel(), _st())
        ctx.act = act
        ctx.tc = tc
        ctx.keys = (W.data_ptr(), None if b is None else b.data_ptr())
        ctx.needs = (ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        ctx.save_for_backward(x, W, z if act == ACT_GELU else (y if act == ACT_SIGMOID else None))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, W, zy = ctx.saved_tensors
        M, K = x.shape
        N = W.shape[0]
        dy = dy.contiguous()
        need_dx, need_dw = ctx.needs
        dz = torch.empty((M, N), dtype=torch.float32, device=dy.device)
        dx = torch.empty((M, K), dtype=torch.float32, device=dy.device) if need_dx else None
        dW = _gout(ctx.keys[0], (N, K), dy.device) if need_dw else None
        db = _gout(ctx.keys[1], (N,), dy.device)
        if ctx.tc:
            ws = _workspace(lib().pb_linear_tc_workspace(M, N, K), dy.device)
            if WGRAD_STREAM and HEAD_WGRAD_SIDE and need_dx and need_dw:
                # dZ, db and dX on the critical path; dW = dZ^T X (reads dZ as it stands) beside it
                _call("pb_linear_bwd_tc", _p(dy), _p(zy), _p(x), _p(W), _p(dz), _p(dx), None, _p(db), M, N, K, ctx.act, 0,
                                             _p(ws), ws.numel(), _st())
                run_on_side(dy.device, lambda: _call("pb_linear_bwd_tc", None, None, _p(x), _p(W), _p(dz), None, _p(dW), None, M, N, K,
                                                     ctx.act, 0, _p(ws), ws.numel(), _st()), x, dz, dW)
            else:
                _call("pb_linear_bwd_tc", _p(dy), _p(zy), _p(x), _p(W), _p(dz), _p(dx), _p(dW), _p(db), M, N, K, ctx.act, 0,
                                             _p(ws), ws.numel(), _st())
        else:
            nbytes = lib().pb_workspace_bytes(b"linear_bwd", M, N, K)
            ws = _workspace(nbytes, dy.device)
            _call("pb_linear_bwd", _p(dy), _p(zy), _p(x), _p(W), _p(dz), _p(dx), _p(dW), _p(db), M, N, K, ctx.act, 0,
                                      _p(ws), ws.numel(), _st())
        return dx, dW, db, None, None


class ResLnRowsFn(torch.autograd.Function):
    """LayerNorm_G(a + b + (sum(w_param)/K) * c)  (modules.py:223-229 and the head mix of :87-92)."""

    @staticmethod
    def forward(ctx, a, b, c, w_param, gamma, beta):
        _need_cuda(a, gamma)
        a = a.contiguous()
        b = None if b is None else b.contiguous()
        c = None if c is None else c.contiguous()
        B, G = a.shape
        K = 0 if w_param is None else w_param.numel()
        xhat = torch.empty_like(a)
        rstd = torch.empty((B,), dtype=torch.float32, device=a.device)
        y = torch.empty_like(a)
        _call("pb_res_ln_rows_fwd", _p(a), _p(b), _p(c), _p(w_param), K, _p(gamma), _p(beta), _p(xhat), _p(rstd), _p(y),
                                       B, G, _st())
        ctx.save_for_backward(gamma, xhat, rstd, c, w_param)
        ctx.has = (b is not None, c is not None)
        ctx.keys = (gamma.data_ptr(), beta.data_ptr(), None if w_param is None else w_param.data_ptr())
        return y

    @staticmethod
    def backward(ctx, dy):
        gamma, xhat, rstd, c, w_param = ctx.saved_tensors
        has_b, has_c = ctx.has
        dy = dy.contiguous()
        B, G = dy.shape
        K = 0 if w_param is None else w_param.numel()
        dx = torch.empty_like(dy)
        dc = torch.empty_like(dy) if has_c else None
        # straight into the gradient bucket when a sink is active (the entry point clears these three outputs itself)
        dgamma = _gout(ctx.keys[0], tuple(gamma.shape), dy.device)
        dbeta = _gout(ctx.keys[1], tuple(gamma.shape), dy.device)
        dwp = _gout(ctx.keys[2], tuple(w_param.shape), dy.device) if has_c else None
        _call("pb_res_ln_rows_bwd", _p(dy), _p(gamma), _p(xhat), _p(rstd), _p(c), _p(w_param), K, _p(dx), _p(dc),
                                       _p(dgamma), _p(dbeta), _p(dwp), B, G, _st())
        return dx, (dx if has_b else None), dc, dwp, dgamma, dbeta


class GlobalLink:
    """Side channel between the two fused global-track functions of one block.  GlobalPostFn.backward produces
    dgh (= dx1) and, because it holds every row anyway, also dgh * gelu'(z3); GlobalPreFn.backward takes that product
    from here instead of recomputing the B x G derivatives in each of its CTAs.  Used only when the tensor autograd
    hands to GlobalPreFn.backward is the very tensor GlobalPostFn.backward returned (no other consumer of gh)."""
    __slots__ = ("z3", "dz3", "dgh_ptr")

    def __init__(self):
        self.z3 = self.dz3 = self.dgh_ptr = None


class GlobalPreFn(torch.autograd.Function):
    """(g2l, gh) = (gelu(g W1^T + b1), gelu(g W3^T + b3)) in one launch (modules.py:208,221)."""

    @staticmethod
    def forward(ctx, g, W1, b1, W3, b3, link=None):
        _need_cuda(g, W1, W3)
        g = g.contiguous()
        B, G = g.shape
        C = W1.shape[0]
        dev = g.device
        zg = torch.empty((2, B, C), dtype=torch.float32, device=dev)
        z3 = torch.empty((2, B, G), dtype=torch.float32, device=dev)
        _call("pb_global_pre_fwd", _p(g), _p(W1), _p(b1), _p(W3), _p(b3), _p(zg[0]), _p(zg[1]), _p(z3[0]), _p(z3[1]),
              B, G, C, _st())
        ctx.save_for_backward(g, W1, W3, zg, z3)
        ctx.link = link
        if link is not None:
            link.z3 = z3[0]
        return zg[1], z3[1]

    @staticmethod
    def backward(ctx, dg2l, dgh):
        g, W1, W3, zg, z3 = ctx.saved_tensors
        B, G = g.shape
        C = W1.shape[0]
        dev = g.device
        dg2l = torch.zeros((B, C), dtype=torch.float32, device=dev) if dg2l is None else dg2l.contiguous()
        dgh = torch.zeros((B, G), dtype=torch.float32, device=dev) if dgh is None else dgh.contiguous()
        link = ctx.link
        ready = 0
        if link is not None and link.dz3 is not None and link.dgh_ptr == dgh.data_ptr():
            dgh, ready = link.dz3, 1
        if link is not None:
            link.z3 = link.dz3 = link.dgh_ptr = None
        dg = torch.empty_like(g)
        dW1, db1 = torch.empty_like(W1), torch.empty((C,), dtype=torch.float32, device=dev)
        dW3, db3 = torch.empty_like(W3), torch.empty((G,), dtype=torch.float32, device=dev)
        if ready and WGRAD_STREAM:
            # the two weight gradients leave the critical path: the kernel publishes dzg, the products run on the side stream
            dzg = torch.empty((B, C), dtype=torch.float32, device=dev)
            _call("pb_global_pre_bwd", _p(dg2l), _p(dgh), _p(zg[0]), _p(z3[0]), _p(g), _p(W1), _p(W3), _p(dg), None, _p(db1),
                  None, _p(db3), B, G, C, ready, _p(dzg), _st())
            run_on_side(dev, lambda: (_call("pb_matmul_tn", _p(dgh), _p(g), _p(dW3), B, G, G, _st()),
                                      _call("pb_matmul_tn", _p(dzg), _p(g), _p(dW1), B, C, G, _st())), dgh, dzg, g, dW3, dW1)
        else:
            _call("pb_global_pre_bwd", _p(dg2l), _p(dgh), _p(zg[0]), _p(z3[0]), _p(g), _p(W1), _p(W3), _p(dg), _p(dW1), _p(db1),
                  _p(dW3), _p(db3), B, G, C, ready, None, _st())
        return dg, dW1, db1, dW3, db3, None


class GlobalPostFn(torch.autograd.Function):
    """g2 = LN(g1 + gelu(g1 W4^T + b4)), g1 = LN(g + gh + (sum(Wp)/K) pool) in one launch (modules.py:92,223-229)."""

    @staticmethod
    def forward(ctx, g, gh, pool, Wp, gamma1, beta1, W4, b4, gamma2, beta2, link=None):
        _need_cuda(g, gh, pool, W4)
        g, gh, pool = g.contiguous(), gh.contiguous(), pool.contiguous()
        B, G = g.shape
        dev = g.device
        buf = torch.empty((5, B, G), dtype=torch.float32, device=dev)      # g1, xhat1, z4, xhat2, g2
        rstd = torch.empty((2, B), dtype=torch.float32, device=dev)
        _call("pb_global_post_fwd", _p(g), _p(gh), _p(pool), _p(Wp), Wp.numel(), _p(gamma1), _p(beta1), _p(W4), _p(b4),
              _p(gamma2), _p(beta2), _p(buf[0]), _p(buf[1]), _p(rstd[0]), _p(buf[2]), _p(buf[3]), _p(rstd[1]), _p(buf[4]),
              B, G, _st())
        ctx.save_for_backward(pool, Wp, gamma1, W4, gamma2, buf, rstd)
        ctx.link = link
        return buf[4]

    @staticmethod
    def backward(ctx, dg2):
        pool, Wp, gamma1, W4, gamma2, buf, rstd = ctx.saved_tensors
        B, G = pool.shape
        dev = pool.device
        dg2 = dg2.contiguous()
        link = ctx.link
        z3 = link.z3 if link is not None else None
        out = torch.empty((6 if z3 is not None else 5, B, G), dtype=torch.float32, device=dev)   # 3 x scratch, dx1, dpool[, dz3]
        dWp = torch.empty_like(Wp)
        vec = torch.empty((5, G), dtype=torch.float32, device=dev)         # dgamma1, dbeta1, db4, dgamma2, dbeta2
        dW4 = torch.empty_like(W4)
        dz3 = out[5] if z3 is not None else None
        side = WGRAD_STREAM
        _call("pb_global_post_bwd", _p(dg2), _p(gamma2), _p(buf[3]), _p(rstd[1]), _p(buf[2]), _p(buf[0]), _p(W4), _p(gamma1),
              _p(buf[1]), _p(rstd[0]), _p(pool), _p(Wp), Wp.numel(), _p(out[0]), _p(out[3]), _p(out[4]), _p(dWp), _p(vec[0]),
              _p(vec[1]), None if side else _p(dW4), _p(vec[2]), _p(vec[3]), _p(vec[4]), _p(z3), _p(dz3), B, G, _st())
        if side:        # dW4 = dz4^T g1 off the critical path (dz4 is the third scratch slab)
            run_on_side(dev, lambda: _call("pb_matmul_tn", _p(out[2]), _p(buf[0]), _p(dW4), B, G, G, _st()), out, buf, dW4)
        if dz3 is not None:
            link.dz3, link.dgh_ptr = dz3, out[3].data_ptr()
        return out[3], out[3], out[4], dWp, vec[0], vec[1], dW4, vec[2], vec[3], vec[4], None


class CutFn(torch.autograd.Function):
    """Identity on a group of tensors, used to mark where GraphedPretrainStep splits backward in two.  torch.autograd.grad
    with respect to non-leaf tensors captures the gradient at the node that PRODUCED them and still executes every node with
    an edge into such a node -- with several producers on the cut (a block's local track and the global-track unit below it)
    one of them would run in both halves.  Behind this node all crossing tensors have ONE producer, and nothing below it
    runs before the second half asks for it."""

    @staticmethod
    def forward(ctx, tag, *ts):
        ctx.tag = tag                     # which cut this is (the index of the block above it)
        return tuple(t.view_as(t) for t in ts)

    @staticmethod
    def backward(ctx, *gs):
        if CUT_HOOK is not None:          # one-graph data-parallel step: mark "everything above this cut has been produced"
            CUT_HOOK(ctx.tag)
        return (None,) + gs


CUT_HOOK = None


def gtrack_ok(B, G, C) -> bool:
    return bool(lib().pb_gtrack_ok(B, G, C))


class GTrackLink:
    """Side channel between consecutive GTrackFn units.  Unit i's pre half produces z3 of block i+1; unit i+1's backward holds
    every row of dv1 (the gradient of gh) anyway and emits dz3 = dv1 * gelu'(z3) beside it, which unit i's backward then
    stages as a plain copy instead of forming it in a row phase of its own.  Used only when the gradient autograd hands to
    unit i for gh is the very tensor unit i+1 returned."""
    __slots__ = ("z3", "dz3", "dgh_ptr")

    def __init__(self):
        self.z3 = self.dz3 = self.dgh_ptr = None


class GTrackFn(torch.autograd.Function):
    """The global track across one block boundary as ONE 16-CTA cluster launch (csrc/gtrack.cu): the post half of block i
    (modules.py:92,223-229: g1 = LN(g + gh + att), g2 = LN(g1 + gelu(g1 W4^T + b4))) followed by the pre half of block i+1
    (:208,:221: g2l = gelu(g2 W1^T + b1), gh = gelu(g2 W3^T + b3)).  W4 is None for the first unit (pre half of block 0
    only), W1 is None for the last (post half only).  Returns (g2, g2l, gh) with None for an absent half."""

    @staticmethod
    def forward(ctx, g, gh, pool, Wp, gamma1, beta1, W4, b4, gamma2, beta2, W1, b1, W3, b3, C, link_in, link_out):
        _need_cuda(g)
        has_post, has_pre = W4 is not None, W1 is not None
        g = g.contiguous()
        B, G = g.shape
        dev = g.device
        buf = rstd = zg2 = z32 = None
        if has_post:
            gh, pool = gh.contiguous(), pool.contiguous()
            buf = torch.empty((5, B, G), dtype=torch.float32, device=dev)          # g1, xhat1, z4, xhat2, g2
            rstd = torch.empty((2, B), dtype=torch.float32, device=dev)
        if has_pre:
            zg2 = torch.empty((2, B, C), dtype=torch.float32, device=dev)          # zg, g2l
            z32 = torch.empty((2, B, G), dtype=torch.float32, device=dev)          # z3, gh
        b = lambda t, i: None if t is None else t[i].data_ptr()
        _call("pb_gtrack_fwd", _p(g) if has_post else None, _p(gh) if has_post else None, _p(pool) if has_post else None,
              _p(Wp) if has_post else None, Wp.numel() if has_post else 0, _p(gamma1), _p(beta1), _p(W4), _p(b4), _p(gamma2), _p(beta2),
              b(buf, 0), b(buf, 1), b(rstd, 0), b(buf, 2), b(buf, 3), b(rstd, 1), b(buf, 4),
              None if has_post else _p(g), _p(W1), _p(b1), _p(W3), _p(b3), b(zg2, 0), b(zg2, 1), b(z32, 0), b(z32, 1), B, G, C, _st())
        ctx.save_for_backward(g, pool, Wp, gamma1, W4, gamma2, buf, rstd, W1, W3, zg2, z32)
        ctx.cfg = (B, G, C, has_post, has_pre)
        ctx.link_in, ctx.link_out = link_in, link_out
        kp = lambda t: None if t is None else t.data_ptr()
        ctx.keys = {n: kp(t) for n, t in (("Wp", Wp), ("gamma1", gamma1), ("beta1", beta1), ("W4", W4), ("b4", b4), ("gamma2", gamma2),
                                          ("beta2", beta2), ("W1", W1), ("b1", b1), ("W3", W3), ("b3", b3))}
        if has_pre and link_out is not None:
            link_out.z3 = z32[0]
        return (buf[4] if has_post else None, zg2[1] if has_pre else None, z32[1] if has_pre else None)

    @staticmethod
    def backward(ctx, dg2, dg2l, dghn):
        g, pool, Wp, gamma1, W4, gamma2, buf, rstd, W1, W3, zg2, z32 = ctx.saved_tensors
        B, G, C, has_post, has_pre = ctx.cfg
        dev = g.device
        link_in, link_out = ctx.link_in, ctx.link_out
        f32 = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)
        dg2 = None if dg2 is None else dg2.contiguous()
        # ---- pre half (block i+1): needs dg2l, dgh of the next block and the direct-path gradient of g2
        dz3_in = dz3_buf = dzg = db1 = db3 = dg_total = None
        if has_pre:
            dg2l = torch.zeros((B, C), dtype=torch.float32, device=dev) if dg2l is None else dg2l.contiguous()
            dghn = torch.zeros((B, G), dtype=torch.float32, device=dev) if dghn is None else dghn.contiguous()
            if link_out is not None and link_out.dz3 is not None and link_out.dgh_ptr == dghn.data_ptr():
                dz3_in = link_out.dz3
            else:
                dz3_buf = f32(B, G)
            if link_out is not None:
                link_out.z3 = link_out.dz3 = link_out.dgh_ptr = None
            dzg, dg_total = f32(B, C), f32(B, G)
        # ---- post half (block i)
        out = dWp = None
        z3_here = link_in.z3 if (has_post and link_in is not None) else None
        if has_post:
            out = f32(6 if z3_here is not None else 5, B, G)                      # dv2, dz4, dg1, dv1, dpool[, dz3 of this block]
            if not has_pre and dg2 is None:
                dg2 = torch.zeros((B, G), dtype=torch.float32, device=dev)
        # the 16-row groups of the batch (one cluster each) meet in the column sums: those outputs are accumulated, zero on
        # entry.  With a gradient sink they are the parameters' bucket slices (cleared once per step by the bucket); otherwise
        # one zero-filled buffer  [dgamma1, dbeta1, db4, dgamma2, dbeta2 | db3 | db1 | dWp]
        Kd = Wp.numel() if has_post else 0
        keys = ctx.keys
        vec_t = None
        sunk = None
        if GRAD_SINK is not None:
            want = ([("gamma1", (G,)), ("beta1", (G,)), ("b4", (G,)), ("gamma2", (G,)), ("beta2", (G,)), ("Wp", (Kd,))] if has_post else []) + \
                   ([("b3", (G,)), ("b1", (C,))] if has_pre else [])
            got = {n: GRAD_SINK.take(keys[n], shp) for n, shp in want}
            if all(t is not None for t in got.values()):
                sunk = got
        if sunk is not None:
            if has_post:
                vec_t = [sunk["gamma1"], sunk["beta1"], sunk["b4"], sunk["gamma2"], sunk["beta2"]]
                dWp = sunk["Wp"]
            if has_pre:
                db3, db1 = sunk["b3"], sunk["b1"]
        else:
            acc = torch.zeros(((5 * G if has_post else 0) + (G + C if has_pre else 0) + Kd,), dtype=torch.float32, device=dev)
            pos = 0
            if has_post:
                vec_t = [acc[i * G:(i + 1) * G] for i in range(5)]
                pos = 5 * G
            if has_pre:
                db3, db1 = acc[pos:pos + G], acc[pos + G:pos + G + C]
                pos += G + C
            if has_post:
                dWp = acc[pos:pos + Kd]
        o = lambda i: None if out is None else out[i].data_ptr()
        v = lambda i: None if vec_t is None else vec_t[i].data_ptr()
        _call("pb_gtrack_bwd", _p(dg2l) if has_pre else None, None if zg2 is None else zg2[0].data_ptr(), _p(dz3_in),
              _p(dghn) if (has_pre and dz3_in is None) else None, (z32[0].data_ptr() if (has_pre and dz3_in is None) else None),
              _p(dg2) if has_pre else None, _p(W1), _p(W3), _p(dz3_buf), _p(dzg), _p(db1), _p(db3), _p(dg_total),
              None if has_pre else _p(dg2), _p(gamma2), (buf[3].data_ptr() if has_post else None), (rstd[1].data_ptr() if has_post else None),
              (buf[2].data_ptr() if has_post else None), _p(W4), _p(gamma1), (buf[1].data_ptr() if has_post else None),
              (rstd[0].data_ptr() if has_post else None), _p(pool) if has_post else None, _p(Wp) if has_post else None,
              Wp.numel() if has_post else 0, _p(z3_here), o(0), o(1), o(2), o(3), (out[5].data_ptr() if z3_here is not None else None),
              o(4), _p(dWp), v(0), v(1), v(2), v(3), v(4), B, G, C, _st())
        if has_post and link_in is not None:
            if z3_here is not None:
                link_in.dz3, link_in.dgh_ptr = out[5], out[3].data_ptr()
        # ---- weight gradients dW = dZ^T X, off the critical path
        dW4 = dW1 = dW3 = None
        jobs = []
        if has_post:
            dW4 = _gout(keys["W4"], tuple(W4.shape), dev)
            jobs.append((out[1], buf[0], dW4, G, G))                               # dz4^T g1
        if has_pre:
            gp = buf[4] if has_post else g                                         # the pre half's input
            dz3 = dz3_in if dz3_in is not None else dz3_buf
            dW3, dW1 = _gout(keys["W3"], tuple(W3.shape), dev), _gout(keys["W1"], tuple(W1.shape), dev)
            jobs.append((dz3, gp, dW3, G, G))
            jobs.append((dzg, gp, dW1, C, G))

        def products():          # one launch for the unit's two or three products
            args = []
            for (a, x, dw, n, k) in jobs + [(None, None, None, 0, 0)] * (3 - len(jobs)):
                args += [_p(a), _p(x), _p(dw), n, k]
            _call("pb_matmul_tn3", *args, B, _st())

        if WGRAD_STREAM:
            run_on_side(dev, products, *[t for j in jobs for t in j[:3]])
        else:
            products()
        if has_post:
            ret_g, ret_gh, ret_pool = out[3], out[3], out[4]          # dv1 is the gradient of g and of gh; dpool = coef * dv1
        else:
            ret_g, ret_gh, ret_pool = dg_total, None, None
        vt = lambda i: None if vec_t is None else vec_t[i]
        return (ret_g, ret_gh, ret_pool, dWp, vt(0), vt(1), dW4, vt(2), vt(3), vt(4), dW1, db1, dW3, db3, None, None, None)


def global_fused_ok(B, G, C) -> bool:
    return bool(lib().pb_global_fused_ok(B, G, C))


# ---- weight gradients off the critical path ------------------------------------------------------------------------
# Only the optimizer reads the weight gradients, so LocalTrackFn.backward can issue the three weight-gradient GEMM groups
# of a block on a second, low-priority stream: they then fill the SMs that the narrow kernels of the critical path leave
# idle (the 64..80-CTA global-track kernels, the 128-CTA convolutions, every kernel's tail).  The side stream is joined
# at the end of each block's backward, or -- inside pretraining_step / GraphedPretrainStep, which set DEFER_WGRAD_JOIN --
# once after the whole backward (join_side_streams()).  PB_B200_WGRAD_STREAM=0 turns the second stream off.
import os as _os

WGRAD_STREAM = _os.environ.get("PB_B200_WGRAD_STREAM", "1") != "0"
HEAD_WGRAD_SIDE = _os.environ.get("PB_B200_HEAD_WGRAD_SIDE", "1") == "1"      # dW of the GO-output layer and of the token head
CONV_WGRAD_EARLY = _os.environ.get("PB_B200_CONV_WGRAD_EARLY", "1") == "1"
DEFER_WGRAD_JOIN = False
_SIDE = {}
_SIDE_PENDING = set()


def _side_stream(dev):
    s = _SIDE.get(dev)
    if s is None:
        s = torch.cuda.Stream(device=dev, priority=0)          # the main stream of a step may be high priority; this is not
        _SIDE[dev] = s
    return s


def run_on_side(dev, fn, *used):
    """Run fn (which launches on the then-current stream) on the side stream, after everything issued so far on the current
    stream; joined at once unless a step function defers the join to the end of backward."""
    side, main = _side_stream(dev), torch.cuda.current_stream(dev)
    side.wait_stream(main)
    with torch.cuda.stream(side):
        fn()
    for t in used:
        if t is not None:
            t.record_stream(side)
    _SIDE_PENDING.add(dev)
    if not DEFER_WGRAD_JOIN:
        join_side_streams()


def join_side_streams():
    """Make the current stream of every device wait for the weight gradients issued on the side stream."""
    for dev in list(_SIDE_PENDING):
        torch.cuda.current_stream(dev).wait_stream(_SIDE[dev])
    _SIDE_PENDING.clear()


class ZeroArena:
    """One zero-filled buffer per model forward from which every block takes its "zero on entry" operands (LayerNorm
    statistics, pooling accumulator, and the accumulated gradients of its backward): one fill launch per step instead of
    three per block.  Slices are handed out once; a LocalTrackFn backward that runs a second time on the same graph
    allocates fresh zeros instead."""

    def __init__(self, nbytes: int, device):
        self.buf = torch.zeros(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        self.off = 0

    @staticmethod
    def _acc_floats(B, L, C):
        return (4 * L * C + 3 * C + B * C + 3) // 4 * 4 + 8 * B

    @staticmethod
    def block_bytes(B, L, C, G):
        def up(n):
            return (n + 255) // 256 * 256
        return up(2 * B * 2 * 8) + up(B * G * 4) + up(ZeroArena._acc_floats(B, L, C) * 4)

    def take(self, shape, dtype):
        n = 1
        for d in shape:
            n *= int(d)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        o = self.off
        if o + nbytes > self.buf.numel():
            return torch.zeros(shape, dtype=dtype, device=self.buf.device)
        self.off = (o + nbytes + 255) // 256 * 256
        return self.buf[o:o + nbytes].view(dtype).view(shape)


class LocalTrackFn(torch.autograd.Function):
    """The local track of one ProteinBERTBlock plus the attention pooling it feeds
    (modules.py:205-219): conv pair + GELU + 4-way add -> joint LN -> Linear+GELU+residual -> joint LN
    -> pool[b,g] = sum_l gelu(h . Wv).  Returns (h2 [B,L,C] act, pool [B,G] f32)."""

    @staticmethod
    def forward(ctx, x, g2l, Wn, bn, Ww, bw, ln1w, ln1b, W2, b2, ln2w, ln2b, staged, ksize, dil, head_grads, arena, *head_wv):
        with_pool = len(head_wv) > 0          # no head tensors: the caller evaluates the attention layer itself (ops.AttentionFn)
        _need_cuda(x, g2l, Wn)
        x = x.contiguous()
        dt = dt_of(x)
        if staged.dt != dt:
            raise _lib.PBError("activation dtype does not match the staged weights' precision")
        B, L, C = x.shape
        G = staged.wvT.shape[0]
        L_ = lib()
        st = _st()
        dev = x.device
        wn_f, wn_b, ww_f, ww_b = staged.wn_f, staged.wn_b, staged.ww_f, staged.ww_b
        W2w, wvTw = staged.W2w, staged.wvT
        # transposed copies: the K-major operands of the tensor-core data-gradient GEMMs
        W2Tw, wvw = staged.W2Tw, staged.wv
        if arena is not None:
            stats = arena.take((2, B, 2), torch.float64)
        else:
            stats = torch.zeros((2, B, 2), dtype=torch.float64, device=dev)
        stats1, stats2 = stats[0], stats[1]
        zn, zw, u1, h1, z2, u2, h2 = (torch.empty_like(x) for _ in range(7))
        g2l = g2l.contiguous()
        _call("pb_conv_pair_fwd", _p(x), _p(wn_f), _p(ww_f), _p(bn), _p(bw), _p(g2l), _p(zn), _p(zw), _p(u1), _p(stats1),
                                  B, L, C, ksize, dil, dt, st)
        _call("pb_ln_joint_apply", _p(u1), _p(stats1), _p(ln1w), _p(ln1b), _p(h1), B, L, C, LN_EPS, dt, st)
        _call("pb_local_linear_fwd", _p(h1), _p(W2w), _p(b2), _p(z2), _p(u2), _p(stats2), B, L, C, dt, st)
        _call("pb_ln_joint_apply", _p(u2), _p(stats2), _p(ln2w), _p(ln2b), _p(h2), B, L, C, LN_EPS, dt, st)
        pool = None
        if with_pool:
            pool = arena.take((B, G), torch.float32) if arena is not None else torch.zeros((B, G), dtype=torch.float32, device=dev)
            _call("pb_attn_pool_fwd", _p(h2), _p(wvTw), _p(pool), B, L, C, G, dt, st)
        ctx.save_for_backward(x, zn, zw, u1, stats1, h1, z2, u2, stats2, h2, ln1w, ln2w)
        # the zeroed accumulators of backward are reserved now, from the same per-step fill
        ctx.bwd_zero = arena.take((ZeroArena._acc_floats(B, L, C),), torch.float32) if arena is not None else None
        ctx.staged = staged
        ctx.cfg = (B, L, C, G, ksize, dil, dt, bool(head_grads), len(head_wv))
        ctx.keys = tuple(t.data_ptr() for t in (Wn, bn, Ww, bw, ln1w, ln1b, W2, b2, ln2w, ln2b))
        return h2, pool

    @staticmethod
    def backward(ctx, dh2, dpool):
        x, zn, zw, u1, stats1, h1, z2, u2, stats2, h2, ln1w, ln2w = ctx.saved_tensors
        sw = ctx.staged
        wn_b, ww_b, W2w, wvTw, W2Tw, wvw = sw.wn_b, sw.ww_b, sw.W2w, sw.wvT, sw.W2Tw, sw.wv
        B, L, C, G, ksize, dil, dt, head_grads, n_heads = ctx.cfg
        L_ = lib()
        st = _st()
        dev = x.device
        adt = act_dtype(dt)
        if dh2 is None:
            dh2 = torch.zeros_like(x)
        dh2 = dh2.contiguous().to(adt)
        if n_heads > 0:
            dpool = torch.zeros((B, G), dtype=torch.float32, device=dev) if dpool is None else dpool.contiguous()

        def f32(*shape):
            return torch.empty(shape, dtype=torch.float32, device=dev)

        # every accumulated ("zero on entry") output of this backward lives in one arena cleared by one fill
        n_acc = (4 * L * C + 3 * C + B * C + 3) // 4 * 4                 # keeps the double area 16-byte aligned
        arena, ctx.bwd_zero = ctx.bwd_zero, None          # a second backward over the same graph must not reuse it
        if arena is None or arena.numel() != n_acc + 8 * B:
            arena = torch.zeros(n_acc + 8 * B, dtype=torch.float32, device=dev)
        dln1w, dln1b = arena[0:L * C].view(L, C), arena[L * C:2 * L * C].view(L, C)
        dln2w, dln2b = arena[2 * L * C:3 * L * C].view(L, C), arena[3 * L * C:4 * L * C].view(L, C)
        o = 4 * L * C
        dbn, dbw, db2 = arena[o:o + C], arena[o + C:o + 2 * C], arena[o + 2 * C:o + 3 * C]
        kWn, kbn, kWw, kbw, kl1w, kl1b, kW2, kb2, kl2w, kl2b = ctx.keys
        if GRAD_SINK is not None:          # accumulated parameter gradients: the bucket's slices (cleared once per step) instead
            got = [GRAD_SINK.take(k_, shp) for k_, shp in ((kl1w, (L, C)), (kl1b, (L, C)), (kl2w, (L, C)), (kl2b, (L, C)),
                                                            (kbn, (C,)), (kbw, (C,)), (kb2, (C,)))]
            if all(t is not None for t in got):
                dln1w, dln1b, dln2w, dln2b, dbn, dbw, db2 = got
        dg2l = arena[o + 3 * C:o + 3 * C + B * C].view(B, C)
        sums_all = arena[n_acc:].view(torch.float64).view(2, B, 2)          # two [B,2] double scratch areas

        side = _side_stream(dev) if WGRAD_STREAM else None
        main = torch.cuda.current_stream(dev)

        def on_side(fn, *used):
            """Run fn (which launches on the then-current stream) on the side stream, after everything issued so far."""
            side.wait_stream(main)
            with torch.cuda.stream(side):
                fn()
            for t in used:
                if t is not None:
                    t.record_stream(side)
            _SIDE_PENDING.add(dev)

        # attention pooling (n_heads == 0: the caller evaluates the attention layer itself, ops.AttentionFn)
        Dv = G // max(n_heads, 1)
        dWv = f32(n_heads, C, Dv) if (head_grads and n_heads > 0) else None          # one contiguous [C, Dv] gradient per head
        if n_heads == 0:
            dh2t, dzv = dh2.clone(), None          # dh2t is reused as scratch below: never autograd's own tensor
        else:
            dzv = torch.empty((B * L, G), dtype=adt, device=dev)
            dh2t = torch.empty_like(x)
            if side is None:
                _call("pb_attn_pool_bwd", _p(h2), _p(wvTw), _p(wvw), _p(dpool), _p(dh2), _p(dzv), _p(dh2t), _p(dWv), Dv, B, L, C, G, dt, *_wgrad_ws(dev), st)
            else:
                _call("pb_attn_pool_bwd", _p(h2), _p(wvTw), _p(wvw), _p(dpool), _p(dh2), _p(dzv), _p(dh2t), None, Dv, B, L, C, G, dt, None, 0, st)
                if dWv is not None:
                    on_side(lambda: _call("pb_attn_pool_bwd", _p(h2), _p(wvTw), _p(wvw), None, None, _p(dzv), None, _p(dWv), Dv, B, L, C, G, dt,
                                          *_wgrad_ws(dev), _st()), h2, dzv, dWv, wvTw, wvw)
        del dzv
        fused = bool(L_.pb_ln_joint_bwd_gelu_ok(B, L, C))
        # LN2 (+ the linear layer's GELU gradient when the fused kernel takes the shape)
        du2 = torch.empty_like(x)
        dW2 = _gout(kW2, (C, C), dev)
        if fused:
            dz2 = torch.empty_like(x)
            _call("pb_ln_joint_bwd_gelu", _p(dh2t), _p(u2), _p(stats2), _p(ln2w), _p(z2), None, _p(du2), _p(dz2), None,
                                          _p(dln2w), _p(dln2b), None, _p(db2), None, _p(sums_all[0]), B, L, C, LN_EPS, dt, st)
            dh1 = dh2t                  # dh2t is dead once LN2's backward has read it
        else:
            _call("pb_ln_joint_bwd", _p(dh2t), _p(u2), _p(stats2), _p(ln2w), _p(du2), _p(dln2w), _p(dln2b), _p(sums_all[0]), B, L, C,
                                     LN_EPS, dt, st)
            dz2 = dh2t                  # reuse as scratch
            dh1 = torch.empty_like(x)
        # local linear
        if side is None:
            _call("pb_local_linear_bwd", _p(du2), None if fused else _p(z2), _p(h1), _p(W2w), _p(W2Tw), _p(dz2), _p(dh1), _p(dW2), _p(db2),
                                         B, L, C, dt, *_wgrad_ws(dev), st)
        else:
            _call("pb_local_linear_bwd", _p(du2), None if fused else _p(z2), _p(h1), _p(W2w), _p(W2Tw), _p(dz2), _p(dh1), None, _p(db2),
                                         B, L, C, dt, None, 0, st)
            on_side(lambda: _call("pb_local_linear_bwd", None, None, _p(h1), _p(W2w), _p(W2Tw), _p(dz2), None, _p(dW2), None,
                                  B, L, C, dt, *_wgrad_ws(dev), _st()), h1, dz2, dW2, W2w, W2Tw)
        # LN1 (+ both convolutions' GELU gradients and the per-sample sums behind dg2l)
        du1 = du2                       # du2 is dead after the linear backward
        dzn = dz2 if side is None else torch.empty_like(x)     # dz2 is still being read by the side stream's dW2 GEMM
        if fused:
            dzw = torch.empty_like(x)
            _call("pb_ln_joint_bwd_gelu", _p(dh1), _p(u1), _p(stats1), _p(ln1w), _p(zn), _p(zw), _p(du1), _p(dzn), _p(dzw),
                                          _p(dln1w), _p(dln1b), _p(dg2l), _p(dbn), _p(dbw), _p(sums_all[1]), B, L, C, LN_EPS, dt, st)
        else:
            _call("pb_ln_joint_bwd", _p(dh1), _p(u1), _p(stats1), _p(ln1w), _p(du1), _p(dln1w), _p(dln1b), _p(sums_all[1]), B, L, C,
                                     LN_EPS, dt, st)
            dzw = dh1                   # dead by now
        # conv pair
        dx = torch.empty_like(x)
        dWn, dWw = _gout(kWn, (C, C, ksize), dev), _gout(kWw, (C, C, ksize), dev)
        if side is None:
            _call("pb_conv_pair_bwd", _p(du1), None if fused else _p(zn), None if fused else _p(zw), _p(x), _p(wn_b), _p(ww_b), _p(dzn),
                                      _p(dzw), _p(dx), _p(dWn), _p(dWw), _p(dbn), _p(dbw), _p(dg2l), B, L, C, ksize, dil, dt,
                                      *_wgrad_ws(dev), st)
        else:
            wgrad = lambda: _call("pb_conv_pair_bwd", None, None, None, _p(x), _p(wn_b), _p(ww_b), _p(dzn), _p(dzw), None, _p(dWn), _p(dWw),
                                  None, None, None, B, L, C, ksize, dil, dt, *_wgrad_ws(dev), _st())
            # with the fused LayerNorm-backward kernel dzn / dzw exist before the data gradient is issued: the weight gradient then
            # depends on that kernel only, not on the data gradient behind it
            early = fused and CONV_WGRAD_EARLY
            if early:
                on_side(wgrad, x, dzn, dzw, dWn, dWw)
            _call("pb_conv_pair_bwd", _p(du1), None if fused else _p(zn), None if fused else _p(zw), _p(x), _p(wn_b), _p(ww_b), _p(dzn),
                                      _p(dzw), _p(dx), None, None, _p(dbn), _p(dbw), _p(dg2l), B, L, C, ksize, dil, dt, None, 0, st)
            if not early:
                on_side(wgrad, x, dzn, dzw, dWn, dWw)
            if not DEFER_WGRAD_JOIN:
                join_side_streams()
        if head_grads and n_heads > 0:
            dheads = tuple(dWv[h] for h in range(n_heads))
        else:
            dheads = (None,) * n_heads
        return (dx, dg2l, dWn, dbn, dWw, dbw, dln1w, dln1b, dW2, db2, dln2w, dln2b, None, None, None, None, None) + dheads


class AttnPoolFn(torch.autograd.Function):
    """pool[b,g] = sum_l gelu(h[b,l,:] . Wv_all[:,g]) as a standalone op (GlobalAttention(Head).forward)."""

    @staticmethod
    def forward(ctx, h, wvT):
        _need_cuda(h, wvT)
        h = h.contiguous()
        dt = dt_of(h)
        B, L, C = h.shape
        G = wvT.shape[0]
        wvTw = to_wt(wvT, dt)
        wvw = to_wt(wvT.t(), dt) if dt == BF16 else None
        pool = torch.zeros((B, G), dtype=torch.float32, device=h.device)
        _call("pb_attn_pool_fwd", _p(h), _p(wvTw), _p(pool), B, L, C, G, dt, _st())
        ctx.save_for_backward(h, wvTw, wvw)
        return pool

    @staticmethod
    def backward(ctx, dpool):
        h, wvTw, wvw = ctx.saved_tensors
        dt = dt_of(h)
        B, L, C = h.shape
        G = wvTw.shape[0]
        dev = h.device
        dzv = torch.empty((B * L, G), dtype=h.dtype, device=dev)
        zero = torch.zeros_like(h)
        dh = torch.empty_like(h)
        dWvT = torch.empty((G, C), dtype=torch.float32, device=dev)
        _call("pb_attn_pool_bwd", _p(h), _p(wvTw), _p(wvw), _p(dpool.contiguous()), _p(zero), _p(dzv), _p(dh), _p(dWvT), 0, B, L, C, G,
                                     dt, *_wgrad_ws(dev), _st())
        return dh, dWvT


class AttentionFn(torch.autograd.Function):
    """GlobalAttention as the reference executes it (modules.py:21-92), one fused launch (csrc/ops_attention.cu): tanh(g Wq),
    tanh(h Wk), scores / sqrt(K), softmax over `axis` (0: the reference's repeated-query axis; 1: sequence positions),
    gelu(h Wv), P.V and the W_parameter mix.  Wq [H,G,K], Wk [H,C,K], Wv [H,C,Dv] are the heads' matrices stacked."""

    @staticmethod
    def forward(ctx, h, g, Wq, Wk, Wv, Wp, axis):
        _need_cuda(h, g, Wq)
        h, g = h.contiguous(), g.contiguous().float()
        Wq, Wk, Wv, Wp = Wq.contiguous().float(), Wk.contiguous().float(), Wv.contiguous().float(), Wp.contiguous().float()
        dt = dt_of(h)
        B, L, C = h.shape
        H, G, K = Wq.shape
        dev = h.device
        f32 = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)
        att, q, p, pool = f32(B, G), f32(B, H, K), f32(B, H, L), f32(B, G)
        _call("pb_attention_fwd", _p(h), _p(g), _p(Wq), _p(Wk), _p(Wv), _p(Wp), _p(att), _p(q), _p(p), _p(pool), B, L, C, G, K, H,
              int(axis), dt, _st())
        ctx.save_for_backward(h, g, Wq, Wk, Wv, Wp, q, p, pool)
        ctx.cfg = (B, L, C, G, K, H, int(axis), dt)
        return att

    @staticmethod
    def backward(ctx, datt):
        h, g, Wq, Wk, Wv, Wp, q, p, pool = ctx.saved_tensors
        B, L, C, G, K, H, axis, dt = ctx.cfg
        dev = h.device
        z = lambda *shape: torch.zeros(shape, dtype=torch.float32, device=dev)
        dh, dg, dWq, dWk, dWv, scratch = z(B, L, C), z(B, G), z(*Wq.shape), z(*Wk.shape), z(*Wv.shape), z(1)
        dWp = torch.empty_like(Wp)
        _call("pb_attention_bwd", _p(h), _p(g), _p(Wq), _p(Wk), _p(Wv), _p(Wp), _p(datt.contiguous().float()), _p(q), _p(p), _p(pool),
              _p(dh), _p(dg), _p(dWq), _p(dWk), _p(dWv), _p(dWp), _p(scratch), B, L, C, G, K, H, axis, dt, _st())
        return dh.to(h.dtype), dg, dWq, dWk, dWv, dWp, None


class TokenHeadFn(torch.autograd.Function):
    """softmax over the batch axis of Linear(C->V)(h)  (modules.py:277-284,304).  bf16 activations with
    tensor-core-friendly shapes take the tcgen05 GEMM path; everything else the one-CTA-per-position kernel."""

    @staticmethod
    def forward(ctx, h, Wo, bo, staged_pad=None):
        _need_cuda(h, Wo, bo)
        h = h.contiguous()
        dt = dt_of(h)
        B, L, C = h.shape
        V = Wo.shape[0]
        dev = h.device
        p = torch.empty((B, L, V), dtype=torch.float32, device=dev)
        tc = dt == BF16 and bool(lib().pb_token_head_tc_ok(B, L, C, V))
        ctx.tc = tc
        ctx.keys = (Wo.data_ptr(), bo.data_ptr())
        if tc:
            # Wo zero-padded to 128 vocabulary rows, and its transpose: the K-major operands of the two GEMMs.  staged_pad:
            # the pair already written by this step's weight-staging launch (ProteinBERT.forward)
            if staged_pad is not None:
                wo_pad, woT_pad = staged_pad
            else:
                pad = torch.zeros((2, 128, C), dtype=torch.bfloat16, device=dev) if C == 128 else None
                wo_pad = pad[0] if pad is not None else torch.zeros((128, C), dtype=torch.bfloat16, device=dev)
                woT_pad = pad[1] if pad is not None else torch.zeros((C, 128), dtype=torch.bfloat16, device=dev)
                wo_pad[:V].copy_(Wo)
                woT_pad[:, :V].copy_(Wo.t())
            logits = torch.empty((B * L, 32), dtype=torch.float32, device=dev)
            _call("pb_token_head_fwd_tc", _p(h), _p(wo_pad), _p(bo), _p(logits), _p(p), B, L, C, V, _st())
            ctx.save_for_backward(h, Wo, p, woT_pad)
        else:
            _call("pb_token_head_fwd", _p(h), _p(Wo.contiguous()), _p(bo), _p(p), B, L, C, V, dt, _st())
            ctx.save_for_backward(h, Wo, p)
        return p

    @staticmethod
    def backward(ctx, dP):
        h, Wo, p = ctx.saved_tensors[:3]
        dt = dt_of(h)
        B, L, C = h.shape
        V = Wo.shape[0]
        dev = h.device
        dP = dP.contiguous().float()
        dh = torch.empty_like(h)
        if ctx.tc:
            woT_pad = ctx.saved_tensors[3]
            dlogit = torch.empty((B * L, 128), dtype=torch.bfloat16, device=dev)
            dWo_pad = torch.empty((128, C), dtype=torch.float32, device=dev)
            dbo = _gout(ctx.keys[1], (V,), dev, zero=True)
            dWo = dWo_pad[:V]
            sl = GRAD_SINK.take(ctx.keys[0], (V, C)) if GRAD_SINK is not None else None

            def wgrad(first=None):          # dWo from the logit gradients (the padded GEMM output cannot alias the bucket)
                _call("pb_token_head_bwd_tc", first, _p(p), _p(h), _p(woT_pad), None, _p(dlogit), None if first is None else _p(dh),
                      _p(dWo_pad), _p(dbo), B, L, C, V, *_wgrad_ws(dev), _st())
                if sl is not None:
                    sl.copy_(dWo)

            if WGRAD_STREAM and HEAD_WGRAD_SIDE:
                _call("pb_token_head_bwd_tc", _p(dP), _p(p), _p(h), _p(woT_pad), None, _p(dlogit), _p(dh), None, _p(dbo),
                      B, L, C, V, None, 0, _st())
                run_on_side(dev, wgrad, p, h, dlogit, dWo_pad, sl)
            else:
                wgrad(_p(dP))
            return dh, (sl if sl is not None else dWo), dbo, None
        dWo = _gout(ctx.keys[0], tuple(Wo.shape), dev)
        dbo = _gout(ctx.keys[1], (V,), dev)
        _call("pb_token_head_bwd", _p(dP), _p(p), _p(h), _p(Wo.contiguous()), None, _p(dh), _p(dWo), _p(dbo), B, L, C, V,
              dt, _st())
        return dh, dWo, dbo, None


class PretrainLossFn(torch.autograd.Function):
    """utils.py:293-294 with CrossEntropyLoss/BCELoss(reduction='none') (dummy_tests.py:132-133), float64."""

    @staticmethod
    def forward(ctx, p_local, p_global, y_local, y_global, w_local, w_global):
        _need_cuda(p_local, p_global, y_local)
        p_local, p_global = p_local.contiguous(), p_global.contiguous()
        B, L, V = p_local.shape
        A = p_global.shape[1]
        y_local = y_local.contiguous()
        y_global = y_global.contiguous().float()
        w_local = w_local.contiguous().double()
        w_global = w_global.contiguous().double()
        loss = torch.empty((3,), dtype=torch.float64, device=p_local.device)
        dPl = torch.empty_like(p_local)
        dPg = torch.empty_like(p_global)
        _call("pb_pretrain_loss", _p(p_local), _p(p_global), _p(y_local), _p(y_global), _p(w_local), _p(w_global),
                                     _p(loss), _p(dPl), _p(dPg), B, L, V, A, _st())
        ctx.save_for_backward(dPl, dPg)
        return loss[0]

    @staticmethod
    def backward(ctx, dloss):
        dPl, dPg = ctx.saved_tensors
        if LOSS_GRAD_IS_ONE:          # pretraining_step without a gradient scale: backward starts from 1, nothing to multiply
            return dPl, dPg, None, None, None, None
        s = dloss.float()
        return dPl * s, dPg * s, None, None, None, None


LOSS_GRAD_IS_ONE = False          # set by training.pretraining_step around loss.backward() when it passes no gradient


def linear_act(x, W, b, act, tc=False):
    return LinearActFn.apply(x, W, b, act, tc)


__all__ = ["ACT_GELU", "ACT_NONE", "ACT_SIGMOID", "BF16", "F32", "EmbeddingFn", "LinearActFn", "ResLnRowsFn",
           "LocalTrackFn", "GlobalPreFn", "GlobalPostFn", "GlobalLink", "CutFn", "AttentionFn", "GTrackFn", "GTrackLink", "gtrack_ok", "ZeroArena", "join_side_streams", "global_fused_ok", "AttnPoolFn", "TokenHeadFn", "PretrainLossFn", "linear_act", "act_dtype", "dt_of", "to_wt", "conv_pack"]
