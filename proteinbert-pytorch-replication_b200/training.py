"""The pretraining step around the model: the loss of the reference loop and data-parallel gradient
averaging.  Host logic only -- the arithmetic is in ops.py / libpb_b200.so.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional

import torch
import torch.distributed as dist

from . import ops


def pretraining_loss(p_local, p_global, Y: Dict[str, torch.Tensor], W: Dict[str, torch.Tensor]) -> torch.Tensor:
    """``mean(CE(p_local.permute(0,2,1), Y.local) * W.local) + mean(BCE(p_global, Y.global) * W.global)``
    exactly as utils.py:293-294 evaluates it with the loss modules of dummy_tests.py:132-133
    (CrossEntropyLoss re-normalises the probabilities over the vocabulary axis; BCELoss clamps logs at -100);
    float64 scalar.  One fused kernel instead of six torch ops; differentiable w.r.t. both inputs."""
    return ops.PretrainLossFn.apply(p_local, p_global, Y["local"], Y["global"], W["local"], W["global"])


def _clear_grads(model):
    """``optimizer.zero_grad(set_to_none=True)`` of utils.py:295, extended to the attention-head tensors: the reference
    keeps them in plain Python lists (modules.py:78-81), so no optimizer ever zeroes or reads their ``.grad`` and it
    would grow by one step's gradient per step; here ``.grad`` always holds the last step's gradient (and the step
    saves four accumulate launches per block)."""
    for p in model.parameters():
        p.grad = None
    heads = getattr(model, "head_tensors", None)
    if heads is not None:
        for t in heads().values():
            t.grad = None


def pretraining_step(model, X, Y, W, reducer: Optional["GradAllReducer"] = None, grad_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
    """forward + loss + backward (utils.py:291-299 without the optimizer).  Gradients are left in ``.grad``
    (averaged over ranks when a reducer is given).  ``grad_scale``: a float64 scalar tensor on the model's device that
    backward starts from instead of 1 (data-parallel steps fold 1/world in here, so the summed gradients are the mean).
    Returns the detached loss."""
    _clear_grads(model)
    p_local, p_global = model(X)
    loss = pretraining_loss(p_local, p_global, Y, W)
    if reducer is not None:
        reducer.prepare()
    ops.DEFER_WGRAD_JOIN = reducer is None          # the eager bucket hooks read gradients as they appear: join per block then
    ops.LOSS_GRAD_IS_ONE = grad_scale is None
    try:
        loss.backward(gradient=grad_scale)
    finally:
        ops.DEFER_WGRAD_JOIN = False
        ops.LOSS_GRAD_IS_ONE = False
        ops.join_side_streams()                     # weight gradients issued on the side stream (ops.LocalTrackFn)
    if reducer is not None:
        reducer.finish()
    return loss.detach()


def _needs_zero(name: str) -> bool:
    """Gradients the kernels ACCUMULATE into (LayerNorm affines, biases, W_parameter): their bucket slices are cleared once per step."""
    return "_norm_" in name or name.endswith(".bias") or name.endswith("W_parameter")


def _is_block_pre_param(name: str) -> bool:
    """A block's global-to-local linear and first global linear (modules.py:208,221: they read the block's INPUT g): the last
    gradients its backward produces."""
    return ".global_to_local_linear_layer." in name or ".global_linear_layer_1." in name


def _is_block_post_param(name: str) -> bool:
    """The post half of a block's global track (modules.py:92,223-229): W_parameter, both global LayerNorms, global linear 2 --
    the first gradients its backward produces (with cluster units: together with the block ABOVE, ops.GTrackFn spans the boundary)."""
    return (name.endswith("global_attention_layer.W_parameter") or ".global_norm_1." in name or ".global_norm_2." in name
            or ".global_linear_layer_2." in name)


class FlatGradBucket:
    """All registered gradients in ONE flat float32 buffer, laid out in the order backward produces them: output heads, blocks
    last to first, then the embedding and the annotation-input layer.  While the step runs with the bucket's sink
    (``ops.GRAD_SINK``), the backward kernels write every gradient straight into its slice and autograd adopts those slices as
    ``.grad``: the data-parallel exchange is then NCCL all-reduces over contiguous ranges of this buffer, with no gather or
    scatter pass and no separate scaling (1/world enters through the loss gradient).  Parameters whose ``.grad`` does not end
    up aliasing its slice (a backward path that does not use the sink) are copied in and out by ``pb_flat_copy``."""

    ALIGN = 128          # floats: every group starts 512-byte aligned (NCCL chunks), every tensor 16-byte aligned

    def __init__(self, model, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        nblk = len(model.proteinBERT_blocks)

        def grp(n):
            if n.startswith("pretraining_"):
                return 0
            if n.startswith("proteinBERT_blocks."):
                return 1 + (nblk - 1 - int(n.split(".")[1]))
            return 1 + nblk

        def rank_in_group(n):            # production order inside a block: post half, local track, pre half; accumulated tensors first in each
            part = 0 if _is_block_post_param(n) else (2 if _is_block_pre_param(n) else 1)
            return 2 * part + (0 if _needs_zero(n) else 1)

        order = sorted(range(len(named)), key=lambda i: (grp(named[i][0]), rank_in_group(named[i][0]), i))
        dev = named[0][1].device
        self.table, self.slices, self.zero_ranges, self.group_ranges = {}, {}, [], []
        self.post_end = {}               # group -> offset where the block's post-half parameters end
        off, cur, cur_rank, gstart = 0, None, None, 0
        for i in order:
            n, p = named[i]
            g, r = grp(n), rank_in_group(n)
            if g != cur:
                if cur is not None:
                    self.group_ranges.append((cur, gstart, off))
                off = (off + self.ALIGN - 1) // self.ALIGN * self.ALIGN
                cur, gstart, cur_rank = g, off, None
            if r != cur_rank:
                if cur_rank is not None and cur_rank < 2 <= r:
                    off = (off + self.ALIGN - 1) // self.ALIGN * self.ALIGN
                    self.post_end[g] = off
                cur_rank = r
            if not p.is_contiguous() or p.dtype != torch.float32:
                raise RuntimeError("FlatGradBucket needs contiguous float32 parameters")
            self.table[p.data_ptr()] = (off, tuple(p.shape))
            self.slices[n] = (off, p.numel())
            if _needs_zero(n):
                if self.zero_ranges and self.zero_ranges[-1][1] == off:
                    self.zero_ranges[-1] = (self.zero_ranges[-1][0], off + (p.numel() + 3) // 4 * 4)
                else:
                    self.zero_ranges.append((off, off + (p.numel() + 3) // 4 * 4))
            off += (p.numel() + 3) // 4 * 4
        self.group_ranges.append((cur, gstart, off))
        self.numel = (off + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        import os
        # the exchange: NCCL's all-reduce, or -- measured faster from 4 GPUs up (8 x B200, 64.7 MB: 157 us against NCCL's 276) --
        # the NVSwitch's in-fabric reduction driven by pb_multimem_allreduce on a bucket in symmetric memory
        want = os.environ.get("PB_B200_EXCHANGE", "auto") if self.world > 1 else "nccl"
        self.exchange, self._symm = "nccl", None
        if want == "nvls" or (want == "auto" and self.world >= 4):
            try:
                import torch.distributed._symmetric_memory as symm
                gname = (group or dist.group.WORLD).group_name
                flat = symm.empty(self.numel, dtype=torch.float32, device=dev)
                flat.zero_()
                hdl = symm.rendezvous(flat, gname)
                if not hdl.multicast_ptr:
                    raise RuntimeError("no multicast address for the symmetric bucket (NVSwitch multicast unavailable)")
                # every rank's copy of the bucket sits behind one multicast address; the mapping covers this buffer on every
                # rank of the group
                self.flat, self._symm, self.exchange = flat, hdl, "nvls"
                # 16 CTAs of 512 threads saturate the exchange on 8 GPUs (8 -> 160 us, 16 -> 156 us, 64 -> 170 us for 64.7 MB)
                self._nvls_ctas = self.nvls_shape(int(os.environ.get("PB_B200_NVLS_CTAS", "16")), int(os.environ.get("PB_B200_NVLS_THREADS", "512")))
                # shape of an exchange that runs BESIDE backward (GraphedPretrainStep's early parts): every outstanding reduction
                # holds a load slot of its SM for the microseconds of the switch round trip, and the compute CTAs sharing that SM
                # queue behind them -- at 16 x 512 x 8 in flight concurrent LayerNorm / reduce kernels ran 3-6x slower and the
                # overlap gained nothing.  8 x B200, step time: no overlap 1930 us, 16x512x8 1921, 8x256x4 1866, 148x32x2 1871,
                # 32x128x2 1843, 16x128x4 1841 (scripts/ddp_sweep.py).
                # ... for the default model's 64.7 MB bucket; a larger bucket has to move more bytes in the same backward, so the CTA
                # count grows with it (BASELINE config 5, 297 MB: 16 CTAs left the exchange 850 us exposed on 2 GPUs, 64 CTAs 345)
                early_ctas = 16 * min(4, max(1, -(-4 * self.numel // (80 << 20))))
                self._nvls_ctas_early = self.nvls_shape(int(os.environ.get("PB_B200_NVLS_EARLY_CTAS", str(early_ctas))),
                                                        int(os.environ.get("PB_B200_NVLS_EARLY_THREADS", "128")),
                                                        int(os.environ.get("PB_B200_NVLS_EARLY_INFLIGHT", "4")))
            except Exception as e:
                if want == "nvls":
                    raise
                import warnings
                warnings.warn(f"NVLS gradient exchange unavailable ({type(e).__name__}: {e}); using NCCL")
        if self.exchange != "nvls":
            self.flat = torch.zeros((self.numel,), dtype=torch.float32, device=dev)
        if self.world > 1:
            lo, hi = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
            self._comm_stream = torch.cuda.Stream(device=dev, priority=min(hi, -1))
        self.params = dict(named)
        self.loose, self._loose_table, self._loose_n = [], None, 0
        self.nblk = nblk
        import numpy as np
        rows = []
        for a, b in self.zero_ranges:            # pieces of <= 64k floats so that the one clearing launch spreads over the GPU
            for s0 in range(a, b, 1 << 16):
                rows.append((0, s0, min(1 << 16, b - s0)))
        arr = np.array(rows, dtype=[("ptr", "<u8"), ("offset", "<i8"), ("numel", "<i8")])
        self._zero_table = torch.from_numpy(arr.view(np.uint8).copy()).to(dev)
        self._zero_n = len(rows)

    @staticmethod
    def nvls_shape(ctas: int, threads: int = 512, in_flight: int = 8) -> int:
        """Launch shape word of pb_multimem_allreduce (include/pb_b200.h)."""
        return (ctas & 0xffff) | ((threads & 0xfff) << 16) | ({8: 0, 1: 1, 2: 2, 4: 3}[in_flight] << 28)

    def sink(self):
        return ops.GradSink(self.flat, self.table)

    def zero(self):
        """Clear the accumulated slices (~1 MB per block) with one launch."""
        if self.flat.is_cuda:
            ops._call("pb_flat_copy", self._zero_table.data_ptr(), self._zero_n, self.flat.data_ptr(), 2, 0.0,
                      torch.cuda.current_stream().cuda_stream)
        else:
            for a, b in self.zero_ranges:
                self.flat[a:b].zero_()

    def view(self, name):
        off, n = self.slices[name]
        return self.flat[off:off + n].view(self.params[name].shape)

    def adopt(self):
        """After a step ran with the sink: note which parameters' .grad really alias their slice; the others ("loose") get
        a pb_flat_copy table so that they still travel with the bucket."""
        import numpy as np
        rows, self.loose = [], []
        base, piece = self.flat.data_ptr(), 1 << 16
        for n, p in self.params.items():
            off, numel = self.slices[n]
            if p.grad is None:
                raise RuntimeError(f"{n} has no gradient after the step")
            if p.grad.data_ptr() == base + 4 * off and p.grad.is_contiguous():
                continue
            if not p.grad.is_contiguous() or p.grad.dtype != torch.float32:
                raise RuntimeError("gradients must be contiguous float32")
            self.loose.append(n)
            for s0 in range(0, numel, piece):
                rows.append((p.grad.data_ptr() + 4 * s0, off + s0, min(piece, numel - s0)))
        self._loose_n = len(rows)
        if rows:
            arr = np.array(rows, dtype=[("ptr", "<u8"), ("offset", "<i8"), ("numel", "<i8")])
            self._loose_table = torch.from_numpy(arr.view(np.uint8).copy()).to(self.flat.device)
        self._ptrs = {n: p.grad.data_ptr() for n, p in self.params.items()}

    def check(self):
        for n, p in self.params.items():
            if p.grad is None or p.grad.data_ptr() != self._ptrs[n]:
                raise RuntimeError(f"gradient tensor of {n} moved since the bucket was built")

    def gather_loose(self):
        if self._loose_n:
            ops._call("pb_flat_copy", self._loose_table.data_ptr(), self._loose_n, self.flat.data_ptr(), 1, 1.0,
                      torch.cuda.current_stream().cuda_stream)

    def scatter_loose(self):
        if self._loose_n:
            ops._call("pb_flat_copy", self._loose_table.data_ptr(), self._loose_n, self.flat.data_ptr(), 0, 1.0,
                      torch.cuda.current_stream().cuda_stream)

    def split_offset(self, cut_block: int, units: bool) -> int:
        """Bucket offset below which everything has been produced once backward has come down to the input of block `cut_block`:
        the heads and the blocks from the top down to it -- and, with cluster units, the post half of the block below, whose
        gradients come out of the unit that spans the boundary."""
        g = 1 + (self.nblk - 1 - cut_block)             # the group of block cut_block; the block below is group g + 1
        if units and (g + 1) in self.post_end:
            return self.post_end[g + 1]
        for (gg, a, b) in self.group_ranges:
            if gg == g:
                return b
        return self.numel

    def all_reduce(self, a: int = 0, b: Optional[int] = None, async_op: bool = False, after_event=None):
        """Sum flat[a:b] over the ranks, in place.  async_op: the exchange runs on the bucket's communication stream, beside the
        current one -- after `after_event` if given (an event recorded inside the running graph), else after everything
        issued so far on the current stream; call .wait() on the returned object to make the current stream wait for it."""
        if self.world <= 1:
            return None
        b = self.numel if b is None else b
        dev = self.flat.device
        cur = torch.cuda.current_stream(dev)
        if not async_op:
            self._exchange(a, b, cur)
            return None
        stream = self._comm_stream
        if after_event is not None:
            for ev in (after_event if isinstance(after_event, (list, tuple)) else [after_event]):
                stream.wait_event(ev)
        else:
            stream.wait_stream(cur)
        with torch.cuda.stream(stream):
            self._exchange(a, b, stream, early=True)
        ev = torch.cuda.Event()
        ev.record(stream)
        flat = self.flat

        class _Work:
            def wait(self_inner):
                torch.cuda.current_stream(dev).wait_event(ev)
                flat.record_stream(torch.cuda.current_stream(dev))
        return _Work()

    def _exchange(self, a, b, stream, early=False):
        """The collective itself, issued with `stream` current.  early: it runs beside backward."""
        if self.exchange == "nvls":
            self._symm.barrier(channel=0)                  # every rank's gradients of this range are complete
            ops._call("pb_multimem_allreduce", self._symm.multicast_ptr, a, b - a, dist.get_rank(self.group), self.world,
                      self._nvls_ctas_early if early else self._nvls_ctas, stream.cuda_stream)
            self._symm.barrier(channel=1)                  # every rank's share is written
        else:
            dist.all_reduce(self.flat[a:b], op=dist.ReduceOp.SUM, group=self.group)


class GraphedPretrainStep:
    """forward + loss + backward captured once in CUDA graphs and replayed per step.

    The step is ~160 short kernel launches; issued one by one from Python they cost more host time than the GPU
    needs to run them, so the public fast path is: copy the batch into static device buffers (one async H2D per
    tensor, from pinned memory if the caller provides it), replay, read the loss.  Gradients are
    written to the parameters' static ``.grad`` tensors, ready for ``optimizer.step()``.  Shapes are fixed at
    construction.

    Data-parallel (torch.distributed initialised with world > 1): every gradient is written by its kernel into ONE flat
    bucket (``FlatGradBucket``; ``.grad`` are views of it, 1/world is folded into the loss gradient).  The step stays ONE
    graph; identity nodes (``ops.CutFn``) at chosen block boundaries record external events inside it, and the part of the
    bucket that is complete above each cut is summed over the ranks on the communication stream while the graph goes on
    below it -- with a small launch shape (few loads in flight per SM), because a full-rate exchange starves the compute
    CTAs it shares SMs with.  What backward produces last (block 0, embedding, annotation-input matrix) is exchanged after
    the graph at full rate.  No gather, scatter or scaling kernels.  ``overlap=False``: one exchange after the graph;
    ``PB_B200_DDP_MODE=split``: two graphs with the first exchange issued between their replays.

        step = GraphedPretrainStep(model, X, Y, W)      # example batch fixes shapes / dtypes
        loss = step(X, Y, W)                            # host or device tensors
    """

    def __init__(self, model, X, Y, W, warmup: int = 3, overlap: bool = True, force_split: bool = False, force_bucket: bool = False):
        import os
        import torch.distributed as dist
        dev = next(model.parameters()).device
        self.model = model
        self.device = dev
        self._staging = self._copy_stream = self._staged_src = None
        self.static = [t.to(dev).clone() for t in (X["local"], X["global"], Y["local"], Y["global"], W["local"], W["global"])]
        sx, sy, sw = self._dicts()
        distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        self.world = dist.get_world_size() if distributed else 1
        nblk = len(model.proteinBERT_blocks)
        # how the exchange overlaps backward (world > 1):  "event": ONE graph; an external event recorded at the cut inside it lets
        # the first all-reduce start on the communication stream while the rest of the graph runs (default);  "split": two
        # graphs with the first all-reduce issued between their replays;  overlap=False: one graph, one all-reduce after it.
        mode = os.environ.get("PB_B200_DDP_MODE", "event") if (distributed and overlap) else "single"
        if force_split:
            mode = "split"
        if nblk < 2:
            mode = "single"
        self.mode = mode
        self.split = mode == "split"
        self.bucket = FlatGradBucket(model) if (distributed or force_bucket or force_split) else None
        self.graph_b = None
        self.launches_outside_graph = 0
        # "event" mode marks several cuts (block indices, top-down): everything above a cut is exchanged while backward goes on
        # below it.  Default: the middle block and block 1 -- what is left for the exposed last exchange is block 0, the
        # embedding and the annotation-input matrix (18.3 of the 64.7 MB; its gradient is the last thing backward produces).
        default_cuts = ([nblk // 2, 1] if nblk >= 4 else [nblk // 2]) if nblk >= 2 else [0]
        cuts = [min(max(int(c), 1), nblk - 1) for c in os.environ.get("PB_B200_DDP_CUT", ",".join(map(str, default_cuts))).split(",")] if nblk >= 2 else [0]
        cuts = sorted(set(cuts), reverse=True)
        if mode != "event":
            cuts = cuts[:1]
        cut = cuts[0]
        units = model.uses_cluster_units(self.static[0].shape[0])
        self.cuts = cuts
        self.cut_offsets = [self.bucket.split_offset(c, units) for c in cuts] if (self.bucket is not None and mode in ("split", "event")) else None
        self.cut_offset = self.cut_offsets[-1] if self.cut_offsets else None          # everything below it goes in the last exchange
        # one pair of events per cut: the capture stream and the weight-gradient side stream
        self._cut_events = {c: (torch.cuda.Event(external=True), torch.cuda.Event(external=True)) for c in cuts} if mode == "event" else {}
        join_at_cut = os.environ.get("PB_B200_DDP_JOIN_AT_CUT", "0") == "1"
        scale = None
        if self.world > 1:
            scale = torch.tensor(1.0 / self.world, dtype=torch.float64, device=dev)      # sum over ranks -> mean
        self._scale = scale

        # ---- two-stage backward: stage A = forward + loss + backward down to the input of block `cut`; stage B = the rest.
        # With cluster units (ops.GTrackFn) the unit that spans the boundary -- post half of block cut-1, pre half of block cut --
        # runs in stage A (its inputs are the cut), so the post-half parameters of block cut-1 are produced above the cut.
        below = f"proteinBERT_blocks.{cut - 1}."
        upper = [p for n, p in model.named_parameters()
                 if n.startswith("pretraining_") or (n.startswith("proteinBERT_blocks.") and int(n.split(".")[1]) >= cut)
                 or (units and n.startswith(below) and _is_block_post_param(n))]
        upper_heads = [h.Wv_parameter for blk in list(model.proteinBERT_blocks)[cut:]
                       for h in blk.global_attention_layer.global_attention_heads if blk.head_grads]

        def with_sink(fn):
            if self.bucket is None:
                return fn()
            ops.GRAD_SINK = self.bucket.sink()
            try:
                return fn()
            finally:
                ops.GRAD_SINK = None

        def stage_a():
            _clear_grads(model)
            if self.bucket is not None:
                self.bucket.zero()
            model._boundary_at = cut
            try:
                p_local, p_global = model(sx)
            finally:
                model._boundary_at = None
            loss = pretraining_loss(p_local, p_global, sy, sw)
            bnd = list(model._boundary)          # the tensors that cross the cut: (h, g) or, with cluster units, (h, g, g2l, gh)
            model._boundary = None
            ops.DEFER_WGRAD_JOIN = True
            try:
                grads = torch.autograd.grad(loss, upper + upper_heads + bnd, grad_outputs=scale, retain_graph=True)
            finally:
                ops.DEFER_WGRAD_JOIN = False
                ops.join_side_streams()
            for p, g in zip(upper + upper_heads, grads):
                p.grad = g
            return loss.detach(), bnd, list(grads[len(grads) - len(bnd):])

        def stage_b(bnd, dbnd):
            ops.DEFER_WGRAD_JOIN = True
            try:
                torch.autograd.backward(bnd, dbnd)
            finally:
                ops.DEFER_WGRAD_JOIN = False
                ops.join_side_streams()

        def cut_hook(tag):           # runs inside backward, when everything above cut `tag` has been issued (ops.CutFn.backward)
            # ... on two streams: the weight gradients are on the side stream, which is NOT joined here (a join would hold the
            # critical path until the side stream has drained); the exchange waits for one event per stream instead
            if join_at_cut:
                ops.join_side_streams()
            ev_main, ev_side = self._cut_events[tag]
            ev_main.record()
            ev_side.record(ops._side_stream(dev))

        def one_graph_step():
            if self.bucket is not None:
                # _clear_grads happens inside pretraining_step; the accumulated bucket slices are cleared first
                self.bucket.zero()
            if mode != "event":
                return pretraining_step(model, sx, sy, sw, grad_scale=scale)
            model._boundary_at, ops.CUT_HOOK = tuple(cuts), cut_hook
            try:
                return pretraining_step(model, sx, sy, sw, grad_scale=scale)
            finally:
                model._boundary_at, ops.CUT_HOOK = None, None
                model._boundary = None

        def full_step():
            if self.split:
                out = with_sink(stage_a)
                with_sink(lambda: stage_b(*out[1:]))
                return out[0]
            return with_sink(one_graph_step)

        # the step is captured on a high-priority stream: when SMs free up, the critical path's CTAs are placed before the
        # weight-gradient CTAs waiting on the (default-priority) side stream.  Measured: 1.761 -> 1.750 ms per step.
        hipri = os.environ.get("PB_B200_CAPTURE_PRIORITY", "1") == "1"
        side = torch.cuda.Stream(device=dev, priority=-1 if hipri else 0)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                full_step()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        _clear_grads(model)
        self.graph = torch.cuda.CUDAGraph()
        cap = {"stream": side} if hipri else {}
        if distributed:
            # NCCL's watchdog thread polls events while this thread captures; under the default (global) capture mode that
            # poll raises "operation not permitted when stream is capturing" in the watchdog and aborts the process
            cap["capture_error_mode"] = "thread_local"
        if self.split:
            with torch.cuda.graph(self.graph, **cap):
                out = with_sink(stage_a)
            self.loss = out[0]
            self.graph_b = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph_b, pool=self.graph.pool(), **cap):
                with_sink(lambda: stage_b(*out[1:]))
            del out
        else:
            with torch.cuda.graph(self.graph, **cap):
                self.loss = with_sink(one_graph_step)
        if self.bucket is not None:         # the captured backward fixed the gradient addresses
            self.bucket.adopt()
            # kernels of this library launched per step outside the graph: the loose-tensor gather/scatter and the NVLS exchanges
            n_ex = (1 + (len(self.cut_offsets) if (self.cut_offsets and not self.bucket.loose) else 0)) if self.bucket.exchange == "nvls" else 0
            self.launches_outside_graph = (2 if self.bucket.loose else 0) + (n_ex if self.world > 1 else 0)

    def _dicts(self):
        s = self.static
        return ({"local": s[0], "global": s[1]}, {"local": s[2], "global": s[3]}, {"local": s[4], "global": s[5]})

    @staticmethod
    def _flat(X, Y, W):
        return (X["local"], X["global"], Y["local"], Y["global"], W["local"], W["global"])

    def prefetch(self, X, Y, W):
        """Start the host->device copy of a FUTURE batch on a side stream, into staging buffers, so that it runs while
        the current step computes (pinned host tensors make it a true async DMA).  The next ``__call__`` with the same
        tensors picks the staged copy up with a device-to-device copy instead of waiting for PCIe."""
        src = self._flat(X, Y, W)
        dev = self.device
        if self._staging is None:
            self._staging = [torch.empty_like(t) for t in self.static]
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._staged_done = torch.cuda.Event()
            self._staging_free = torch.cuda.Event()
            self._staging_free.record(torch.cuda.current_stream(dev))
        self._copy_stream.wait_event(self._staging_free)         # the previous staged batch has been moved out
        with torch.cuda.stream(self._copy_stream):
            for dst, t in zip(self._staging, src):
                dst.copy_(t, non_blocking=True)
            self._staged_done.record(self._copy_stream)
        self._staged_src = src          # strong references: an id() could be reused by a new tensor once these are freed

    def _load(self, X, Y, W):
        src = self._flat(X, Y, W)
        cur = torch.cuda.current_stream(self.device)
        staged = self._staged_src
        if staged is not None and len(staged) == len(src) and all(a is b for a, b in zip(staged, src)):
            cur.wait_event(self._staged_done)
            for dst, t in zip(self.static, self._staging):
                dst.copy_(t, non_blocking=True)
            self._staging_free.record(cur)
            self._staged_src = None
        else:
            for dst, t in zip(self.static, src):
                if t is not dst:
                    dst.copy_(t, non_blocking=True)

    def read_loss_async(self):
        """Start copying the loss of the step just issued to pinned host memory and return a handle; ``handle.value()`` waits
        for THAT copy only.  A loop that calls ``.item()`` on the returned loss tensor every step (utils.py:295 does) stalls the
        host until the step has finished and leaves the GPU idle until the next replay is issued; reading the handle one step
        late keeps one step queued behind the running one."""
        if getattr(self, "_loss_host", None) is None:
            self._loss_host = [torch.empty((), dtype=torch.float64).pin_memory() for _ in range(2)]
            self._loss_ev = [torch.cuda.Event(), torch.cuda.Event()]
            self._loss_i = 0
        i = self._loss_i
        self._loss_i ^= 1
        host, ev = self._loss_host[i], self._loss_ev[i]
        host.copy_(self.loss, non_blocking=True)
        ev.record(torch.cuda.current_stream(self.device))

        class _Handle:
            def value(self_inner) -> float:
                ev.synchronize()
                return float(host)
        return _Handle()

    def replay_local(self):
        """The captured graphs alone: this rank's gradients (already scaled by 1/world), no exchange."""
        self.graph.replay()
        if self.graph_b is not None:
            self.graph_b.replay()

    def __call__(self, X, Y, W, prefetch=None):
        """One step on the batch (X, Y, W).  ``prefetch=(Xn, Yn, Wn)`` starts copying the NEXT batch while this one runs."""
        self._load(X, Y, W)
        self.graph.replay()
        if prefetch is not None:
            self.prefetch(*prefetch)
        bk = self.bucket
        if bk is None or bk.world == 1:
            if self.graph_b is not None:
                self.graph_b.replay()
            return self.loss
        work = self._exchange_first()                  # NCCL on its own stream ...
        if self.graph_b is not None:
            self.graph_b.replay()                      # ... while the second half of backward runs
        self._exchange_rest(work)
        return self.loss

    def _exchange_first(self):
        """Start the all-reduce of what backward has produced above the cut (the bucket up to cut_offset): after graph A in
        "split" mode, after the event recorded at the cut inside the running graph in "event" mode.  None when there is
        nothing to start early."""
        bk = self.bucket
        if bk.loose:
            return None
        if self.mode == "split":
            return bk.all_reduce(0, self.cut_offset, async_op=True)
        if self.mode == "event":
            work, a = None, 0
            for c, off in zip(self.cuts, self.cut_offsets):          # top-down; the exchanges queue on the communication stream
                if off > a:
                    work = bk.all_reduce(a, off, async_op=True, after_event=list(self._cut_events[c]))
                    a = off
            return work
        return None

    def _exchange_rest(self, work):
        bk = self.bucket
        bk.gather_loose()
        if work is None:
            bk.all_reduce()
        else:
            bk.all_reduce(self.cut_offset, None)
            work.wait()
        bk.scatter_loose()

    # ---- data-parallel diagnostics (bench.py, scripts/ddp_check.py) ---------------------------------------------------------
    def ddp_check(self, X, Y, W) -> float:
        """max over parameter groups of |grad after the exchange - sum over ranks of the rank-local gradients| / |that sum|.
        The rank-local gradients carry the 1/world factor already, so their sum is the mean the reference recipe
        ("the reference per shard, gradients averaged") asks for."""
        bk = self.bucket
        if bk is None:
            return 0.0
        self._load(X, Y, W)
        self.replay_local()
        bk.gather_loose()
        local = bk.flat.clone()
        if bk.world > 1:
            parts = [torch.empty_like(local) for _ in range(bk.world)]
            dist.all_gather(parts, local, group=bk.group)
            expect = torch.stack(parts).double().sum(0)
            # the very exchange __call__ performs (same ranges, same order), on the gradients that were just gathered:
            # a second replay would differ from the first by the order of the kernels' float atomics
            self._exchange_rest(self._exchange_first())
        else:
            expect = local.double()
        torch.cuda.synchronize(self.device)
        bk.gather_loose()
        worst = 0.0
        for (_, a, b) in bk.group_ranges:
            den = float(expect[a:b].norm())
            err = float((bk.flat[a:b].double() - expect[a:b]).norm())
            worst = max(worst, err / den if den > 0 else err)
        return worst

    def comm_profile(self, X, Y, W, steps: int = 10):
        """Per step, CUDA-event timed, max over ranks: the graphs alone, the whole step, and the all-reduce of the bucket alone."""
        bk = self.bucket
        if bk is None or bk.world == 1:
            return {}
        dev = self.device

        def timed(fn):
            for _ in range(2):
                fn()
            dist.barrier()
            torch.cuda.synchronize(dev)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(steps):
                fn()
            b.record()
            torch.cuda.synchronize(dev)
            t = torch.tensor([a.elapsed_time(b) / steps * 1e3], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t[0])

        self._load(X, Y, W)
        local_us = timed(self.replay_local)
        full_us = timed(lambda: self(X, Y, W))
        ar_us = timed(lambda: bk.all_reduce())
        return {"exposed_comm_us": full_us - local_us, "step_us": full_us, "compute_only_us": local_us,
                "allreduce_alone_us": ar_us, "allreduce_bytes": 4 * bk.numel, "bucket_loose_tensors": len(bk.loose)}


class GradAllReducer:
    """Data-parallel gradient averaging overlapped with backward.

    The reference is single-process; batch sharding is the natural partition of this path, and its one
    exchange step is the all-reduce of the 133 registered gradients (the attention-head tensors are not
    registered in the reference, so -- as there -- they are neither broadcast nor reduced).

    Parameters are grouped into buckets in REVERSE registration order (roughly the order backward produces
    them: the annotation-output layer first, the annotation-input layer last).  A post-accumulate hook on
    each parameter counts its bucket down; when a bucket is complete its gradients are packed into one flat
    buffer and all-reduced asynchronously (NCCL runs it on its own stream over NVLink/NVSwitch) while
    backward continues.  ``finish()`` waits for the outstanding buckets, scales by 1/world and unpacks.
    Works with any backend (``gloo`` in the CPU tests)."""

    def __init__(self, params: Iterable[torch.nn.Parameter], bucket_bytes: int = 8 << 20, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        self.buckets: List[List[torch.nn.Parameter]] = []
        cur, size = [], 0
        for p in reversed(self.params):
            cur.append(p)
            size += p.numel() * p.element_size()
            if size >= bucket_bytes:
                self.buckets.append(cur)
                cur, size = [], 0
        if cur:
            self.buckets.append(cur)
        self._bucket_of = {id(p): i for i, b in enumerate(self.buckets) for p in b}
        self._flat: List[Optional[torch.Tensor]] = [None] * len(self.buckets)
        self._pending = [0] * len(self.buckets)
        self._work = []
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]
        self._armed = False

    def broadcast_parameters(self, src: int = 0, extra_tensors=None):
        """Make every rank start from rank ``src``'s weights.  ``extra_tensors``: tensors that take part in the forward
        pass without being registered parameters -- pass ``model.head_tensors().values()``: the reference keeps the
        attention heads' Wq/Wk/Wv in a plain list (modules.py:73-81), so ``model.parameters()`` does not reach them, yet
        Wv enters every block's pooling.  Ranks that are not seeded identically would otherwise train different models
        and average their gradients without any error."""
        if self.world > 1:
            for p in self.params:
                dist.broadcast(p.data, src, group=self.group)
            for t in (extra_tensors or ()):
                dist.broadcast(t.data, src, group=self.group)

    def prepare(self):
        self._pending = [len(b) for b in self.buckets]
        self._work = []
        self._armed = True

    def _on_grad(self, p):
        if not self._armed or self.world == 1:
            return
        i = self._bucket_of[id(p)]
        self._pending[i] -= 1
        if self._pending[i] == 0:
            self._launch(i)

    def _launch(self, i):
        bucket = self.buckets[i]
        flat = torch.cat([p.grad.reshape(-1) for p in bucket])
        self._flat[i] = flat
        self._work.append((i, dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)))

    def finish(self):
        if not self._armed:
            return
        self._armed = False
        if self.world == 1:
            return
        # parameters that received no gradient this step still have to take part in the collective
        for i, n in enumerate(self._pending):
            if n > 0:
                for p in self.buckets[i]:
                    if p.grad is None:
                        p.grad = torch.zeros_like(p)
                self._pending[i] = 0
                self._launch(i)
        for i, work in self._work:
            work.wait()
            flat = self._flat[i]
            flat.mul_(1.0 / self.world)
            off = 0
            for p in self.buckets[i]:
                n = p.numel()
                p.grad.copy_(flat[off:off + n].view_as(p))
                off += n
            self._flat[i] = None
        self._work = []

    def remove(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []


class FusedAdam(torch.optim.Adam):
    """``torch.optim.Adam`` (the optimizer of dummy_tests.py:127-130 / utils.py:301) whose ``step()`` is ONE kernel launch over
    all parameters (``pb_adam_step``).  Same constructor, ``param_groups``, ``state`` keys (``step``, ``exp_avg``,
    ``exp_avg_sq``) and ``state_dict()`` as the torch class, so checkpoints are interchangeable; ``amsgrad``, ``maximize``,
    ``capturable`` and ``differentiable`` are refused.  Parameters and gradients must be contiguous float32 CUDA tensors.
    The step is HBM-bound: 7 x 4 bytes per parameter (16.18 M parameters: 453 MB, 69 us at the measured copy bandwidth)."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, foreach=False, fused=False)
        self._tables = {}

    @torch.no_grad()
    def step(self, closure=None):
        import numpy as np
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            if group.get("amsgrad") or group.get("maximize") or group.get("capturable") or group.get("differentiable"):
                raise _lib_error("FusedAdam supports plain Adam only (no amsgrad / maximize / capturable / differentiable)")
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            # the device job table holds raw pointers: it is valid only while every parameter, gradient and moment tensor
            # stays where it was (gradients are re-allocated by eager autograd, static under a graph; load_state_dict
            # replaces the moments)
            gkey = tuple((p.data_ptr(), p.grad.data_ptr(),
                          self.state[p]["exp_avg"].data_ptr() if "exp_avg" in self.state[p] else 0,
                          self.state[p]["exp_avg_sq"].data_ptr() if "exp_avg_sq" in self.state[p] else 0) for p in ps)
            tab = self._tables.get(gi)
            if tab is None or tab["gkey"] != gkey or tab["n"] != len(ps):
                self._sync_step_counters()                        # the counters are read below
                for p in ps:
                    if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and p.grad.dtype == torch.float32
                            and p.grad.is_contiguous()):
                        raise _lib_error("FusedAdam needs contiguous float32 CUDA parameters and gradients")
                    st = self.state[p]
                    if len(st) == 0:
                        st["step"] = torch.tensor(0.0, dtype=torch.float32)
                        st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                        st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                steps = {int(self.state[p]["step"]) for p in ps}
                if len(steps) != 1:
                    raise _lib_error("FusedAdam: parameters of one group are at different step counts")
                rows, piece = [], 1 << 16
                for p in ps:
                    m, v = self.state[p]["exp_avg"], self.state[p]["exp_avg_sq"]
                    for s0 in range(0, p.numel(), piece):
                        rows.append((p.data_ptr() + 4 * s0, p.grad.data_ptr() + 4 * s0, m.data_ptr() + 4 * s0, v.data_ptr() + 4 * s0,
                                     min(piece, p.numel() - s0)))
                arr = np.array(rows, dtype=[("p", "<u8"), ("g", "<u8"), ("m", "<u8"), ("v", "<u8"), ("numel", "<i8")])
                gkey = tuple((p.data_ptr(), p.grad.data_ptr(), self.state[p]["exp_avg"].data_ptr(),
                              self.state[p]["exp_avg_sq"].data_ptr()) for p in ps)
                tab = {"gkey": gkey, "n": len(ps), "table": torch.from_numpy(arr.view(np.uint8).copy()).to(ps[0].device),
                       "rows": len(rows), "t": steps.pop(), "steps": [self.state[p]["step"] for p in ps],
                       "stream_dev": ps[0].device}
                self._tables[gi] = tab
            tab["t"] += 1
            b1, b2 = group["betas"]
            ops._call("pb_adam_step", tab["table"].data_ptr(), tab["rows"], float(group["lr"]), float(b1), float(b2), float(group["eps"]),
                      float(group["weight_decay"]), tab["t"], torch.cuda.current_stream(tab["stream_dev"]).cuda_stream)
        return loss

    def _sync_step_counters(self):
        # torch's state_dict carries one step counter per parameter; they are brought up to date only when somebody looks
        for tab in self._tables.values():
            for st in tab["steps"]:
                st.fill_(float(tab["t"]))

    def state_dict(self):
        self._sync_step_counters()
        return super().state_dict()

    def load_state_dict(self, state_dict):
        """torch replaces the ``exp_avg`` / ``exp_avg_sq`` / ``step`` tensors here: the cached device tables (raw pointers
        to the old moments and the old step count) must not survive it."""
        self._tables = {}
        super().load_state_dict(state_dict)
        for st in self.state.values():          # torch moves `step` to the parameter's device for non-capturable Adam or not,
            if "step" in st and torch.is_tensor(st["step"]):     # depending on version: keep it a CPU scalar tensor as step() expects
                st["step"] = st["step"].detach().to("cpu", torch.float32)

    def add_param_group(self, param_group):
        if hasattr(self, "_tables"):
            self._tables = {}
        super().add_param_group(param_group)


class WarmupThenPlateau:
    """The learning-rate schedule utils.pretrain intends (utils.py:253-264): linear warm-up ``lr * step / warmup_duration`` for
    ``warmup_duration`` steps (so the very first step runs at lr = 0, as there), then ``ReduceLROnPlateau(mode='min',
    patience=...)`` on the training loss.  The reference builds it as ``SequentialLR([LambdaLR, ReduceLROnPlateau])``, which
    torch >= 2.2 refuses (ReduceLROnPlateau is not an LRScheduler there) and which never passed the loss to
    the plateau scheduler; this object drives the same two torch schedulers directly.  ``state_dict()`` carries the three
    entries the reference checkpoint stores."""

    def __init__(self, optimizer, warmup_duration: int, patience: int):
        from torch.optim.lr_scheduler import LambdaLR, ReduceLROnPlateau
        self.warmup_duration = max(int(warmup_duration), 1)
        self.plateau = ReduceLROnPlateau(optimizer=optimizer, mode="min", patience=patience)
        self.warmup = LambdaLR(optimizer, lr_lambda=lambda step: float(step / self.warmup_duration))
        self.optimizer = optimizer
        self.steps = 0

    def get_last_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def step(self, loss=None):
        self.steps += 1
        if self.steps < self.warmup_duration:
            self.warmup.step()
        elif self.steps == self.warmup_duration:
            self.warmup.step()                              # reaches the base learning rate
        elif loss is not None:
            self.plateau.step(loss)

    def state_dict(self):
        wsd = {k: v for k, v in self.warmup.state_dict().items() if k != "lr_lambdas"}      # the lambda is rebuilt, not pickled
        return {"scheduler_state_dict": self.plateau.state_dict(), "warmup_scheduler_state_dict": wsd,
                "full_scheduler_state_dict": {"steps": self.steps, "warmup_duration": self.warmup_duration}}

    def load_state_dict(self, sd):
        self.plateau.load_state_dict(sd["scheduler_state_dict"])
        w = dict(sd["warmup_scheduler_state_dict"])
        w.pop("lr_lambdas", None)
        self.warmup.__dict__.update(w)
        self.steps = int(sd["full_scheduler_state_dict"]["steps"])


def pretrain(model, train_dataloader, optimizer, local_loss_fn=None, global_loss_fn=None, max_batch_iterations: int = 1000,
             save_path: str = ".", nb_iterations_checkpoint: int = 1000, optim_scheduler_patience: int = 25,
             warmup_duration: int = 10000, loaded_checkpoint=None, device=None, use_graph: bool = True, log_every: int = 1):
    """``utils.pretrain`` (utils.py:220-345) on the kernels of this package: same arguments, same results dictionary, same
    checkpoint dictionary (``current_batch_iteration``, ``model_state_dict``, ``optimizer_state_dict``, three scheduler
    entries, ``loss``) and the same whole-model pickle at the end.  Differences, all in the shell and none in the arithmetic:

    * the scheduler is WarmupThenPlateau (the reference's SequentialLR construction does not start on torch >= 2.2);
    * forward + loss + backward run as GraphedPretrainStep once the batch shape repeats (``use_graph``); batches of another
      shape (the last one of an epoch) take the eager step.  ``local_loss_fn`` / ``global_loss_fn`` are accepted for
      signature compatibility; the loss is the fused kernel's, equal to ``CrossEntropyLoss(reduction='none')`` /
      ``BCELoss(reduction='none')`` in the reference expression (utils.py:293-294);
    * the loss is not read with ``.item()`` every iteration (a device synchronisation per step, utils.py:295): it is copied
      to pinned host memory asynchronously and logged / fed to the plateau scheduler one iteration late."""
    import logging
    from datetime import datetime
    from pathlib import Path
    from timeit import default_timer as time

    dev = torch.device(device) if device is not None else next(model.parameters()).device
    results = {"train_loss": []}
    sched = WarmupThenPlateau(optimizer, warmup_duration, optim_scheduler_patience)
    if loaded_checkpoint is not None:
        logging.info("Loading checkpoint...")
        current = loaded_checkpoint["current_batch_iteration"]
        model.load_state_dict(loaded_checkpoint["model_state_dict"])
        optimizer.load_state_dict(loaded_checkpoint["optimizer_state_dict"])
        sched.load_state_dict(loaded_checkpoint)
        logging.info("Checkpoint loaded!")
    else:
        current = 0
    model.train()
    graphs = {}                       # batch shape -> GraphedPretrainStep
    pending = []                      # (iteration, pinned host scalar, event, start time)
    last_loss = None

    def drain(block: bool):
        nonlocal last_loss
        while pending and (block or pending[0][2].query()):
            it, host, ev, t0 = pending.pop(0)
            ev.synchronize()
            last_loss = float(host[0])
            results["train_loss"].append(last_loss)
            if log_every and it % log_every == 0:
                logging.info(f"Current batch iteration: {it} | Train loss: {last_loss:.4f} | "
                             f"Learning rate: {sched.get_last_lr()[0]} | Batch Iteration time: {time() - t0:.4f} seconds")

    while current < max_batch_iterations:
        n_before = current
        for X, Y, W in train_dataloader:
            t0 = time()
            X = {k: v.to(dev, non_blocking=True) for k, v in X.items()}
            Y = {k: v.to(dev, non_blocking=True) for k, v in Y.items()}
            W = {k: v.to(dev, non_blocking=True) for k, v in W.items()}
            key = (tuple(X["local"].shape), tuple(X["global"].shape))
            if use_graph and key in graphs:
                loss = graphs[key](X, Y, W)
            else:
                loss = pretraining_step(model, X, Y, W)
                if use_graph and key not in graphs and graphs.get("seen") == key:
                    graphs[key] = GraphedPretrainStep(model, X, Y, W, warmup=1)      # the shape repeated: capture it
                graphs["seen"] = key
            optimizer.step()
            host = torch.empty(1, dtype=torch.float64).pin_memory()
            host.copy_(loss.detach().reshape(1), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            current += 1
            pending.append((current, host, ev, t0))
            drain(block=len(pending) > 2)
            sched.step(last_loss)
            if current >= max_batch_iterations:
                break
            if current % nb_iterations_checkpoint == 0:
                drain(block=True)
                logging.info(f"Saving checkpoint to {save_path}...")
                ck = {"current_batch_iteration": current, "model_state_dict": model.state_dict(),
                      "optimizer_state_dict": optimizer.state_dict(), "loss": results["train_loss"][-1]}
                ck.update(sched.state_dict())
                torch.save(ck, Path(save_path) / f"proteinbert_pretraining_checkpoint_{current}.pt")
        if current == n_before:
            raise RuntimeError("pretrain: the data loader yielded no batches")
    drain(block=True)
    graphs.clear()
    stamp = datetime.now().strftime('%m-%d-%Y_%H-%M-%S')
    logging.info(f"Save whole model to {save_path}...")
    torch.save(model, Path(save_path) / f"proteinbert_pretrained_model_{stamp}.pt")
    return results


def _lib_error(msg):
    from ._lib import PBError
    return PBError(msg)
