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


def pretraining_step(model, X, Y, W, reducer: Optional["GradAllReducer"] = None) -> torch.Tensor:
    """forward + loss + backward (utils.py:291-299 without the optimizer).  Gradients are left in ``.grad``
    (averaged over ranks when a reducer is given).  Returns the detached loss."""
    _clear_grads(model)
    p_local, p_global = model(X)
    loss = pretraining_loss(p_local, p_global, Y, W)
    if reducer is not None:
        reducer.prepare()
    ops.DEFER_WGRAD_JOIN = reducer is None          # the eager bucket hooks read gradients as they appear: join per block then
    try:
        loss.backward()
    finally:
        ops.DEFER_WGRAD_JOIN = False
        ops.join_side_streams()                     # weight gradients issued on the side stream (ops.LocalTrackFn)
    if reducer is not None:
        reducer.finish()
    return loss.detach()


class FlatGradBucket:
    """All registered gradients as ONE flat buffer: a gather kernel after backward, one NCCL all-reduce over
    NVLink/NVSwitch, a scatter kernel that also applies 1/world.  Used by GraphedPretrainStep when the job is
    data-parallel (the gradient tensors have fixed addresses once the step is captured)."""

    def __init__(self, params, group=None):
        import numpy as np
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.params = [p for p in params if p.requires_grad]
        if any(p.grad is None for p in self.params):
            raise RuntimeError("FlatGradBucket needs materialised .grad tensors (run one step first)")
        dev = self.params[0].device
        rows, off = [], 0
        for p in self.params:
            g = p.grad
            if not g.is_contiguous() or g.dtype != torch.float32:
                raise RuntimeError("gradients must be contiguous float32")
            # big tensors are cut into 64k-element pieces so the one-launch copy spreads over the whole GPU
            piece = 1 << 16
            for s0 in range(0, g.numel(), piece):
                rows.append((g.data_ptr() + 4 * s0, off + s0, min(piece, g.numel() - s0)))
            off += (g.numel() + 3) // 4 * 4                      # keep every tensor 16-byte aligned in the bucket
        self.numel = off
        self.flat = torch.zeros((off,), dtype=torch.float32, device=dev)
        arr = np.array(rows, dtype=[("ptr", "<u8"), ("offset", "<i8"), ("numel", "<i8")])
        self.table = torch.from_numpy(arr.view(np.uint8).copy()).to(dev)
        self.n = len(rows)
        self._ptrs = [p.grad.data_ptr() for p in self.params]

    def check(self):
        if [p.grad.data_ptr() for p in self.params] != self._ptrs:
            raise RuntimeError("gradient tensors moved since the bucket was built")

    def gather(self):
        ops._call("pb_flat_copy", self.table.data_ptr(), self.n, self.flat.data_ptr(), 1, 1.0,
                  torch.cuda.current_stream().cuda_stream)

    def all_reduce(self, async_op: bool = False):
        if self.world > 1:
            return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
        return None

    def scatter(self):
        ops._call("pb_flat_copy", self.table.data_ptr(), self.n, self.flat.data_ptr(), 0, 1.0 / self.world,
                  torch.cuda.current_stream().cuda_stream)


class GraphedPretrainStep:
    """forward + loss + backward captured once in a CUDA graph and replayed per step.

    The step is ~170 short kernel launches; issued one by one from Python they cost more host time than the GPU
    needs to run them, so the public fast path is: copy the batch into static device buffers (one async H2D per
    tensor, from pinned memory if the caller provides it), replay the graph, read the loss.  Gradients are
    written to the parameters' static ``.grad`` tensors, ready for ``optimizer.step()``.  Shapes are fixed at
    construction.  Under torch.distributed (world > 1) the registered gradients are then gathered into one flat bucket
    (one kernel), all-reduced by NCCL over NVLink/NVSwitch and scattered back with 1/world applied (one kernel).
    ``overlap=True`` captures backward as TWO graphs split at the middle block instead and all-reduces the first
    half's bucket asynchronously while the second graph replays.  Measured on 2 x B200 (2.08 ms single-GPU step):
    2.31 ms without, 2.39 ms with the split -- the NCCL kernel takes SMs from the one-wave persistent kernels of the
    second graph, which costs more than the 0.07 ms of all-reduce it hides -- so the default is off.

        step = GraphedPretrainStep(model, X, Y, W)      # example batch fixes shapes / dtypes
        loss = step(X, Y, W)                            # host or device tensors
    """

    def __init__(self, model, X, Y, W, warmup: int = 3, overlap: bool = False, force_split: bool = False):
        dev = next(model.parameters()).device
        self.model = model
        self.device = dev
        self._staging = self._copy_stream = self._staged_src = None
        self.static = [t.to(dev).clone() for t in (X["local"], X["global"], Y["local"], Y["global"], W["local"], W["global"])]
        sx, sy, sw = self._dicts()
        distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        nblk = len(model.proteinBERT_blocks)
        self.split = ((distributed and overlap) or force_split) and nblk >= 2
        self.bucket = self.bucket_a = self.bucket_b = None
        self.graph_b = None
        cut = nblk // 2

        # ---- two-stage backward (data-parallel): stage A = forward + loss + backward down to the input of block
        # `cut`; stage B = the rest.  The gradients of stage A are all-reduced while stage B runs.
        # with cluster units (ops.GTrackFn) the pre half of block `cut` (its global-to-local and first global linear) is
        # computed -- and differentiated -- together with the post half of block cut-1, i.e. below the cut
        straddle = (f"proteinBERT_blocks.{cut}.global_to_local_linear_layer.", f"proteinBERT_blocks.{cut}.global_linear_layer_1.") \
            if model.uses_cluster_units(self.static[0].shape[0]) else ()
        upper = [p for n, p in model.named_parameters()
                 if n.startswith("pretraining_") or (n.startswith("proteinBERT_blocks.") and int(n.split(".")[1]) >= cut
                                                     and not n.startswith(straddle))]
        upper_ids = {id(p) for p in upper}
        lower = [p for p in model.parameters() if id(p) not in upper_ids]
        upper_heads = [h.Wv_parameter for blk in list(model.proteinBERT_blocks)[cut:]
                       for h in blk.global_attention_layer.global_attention_heads if blk.head_grads]

        def stage_a():
            _clear_grads(model)
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
                grads = torch.autograd.grad(loss, upper + upper_heads + bnd, retain_graph=True)
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

        def full_step():
            if self.split:
                out = stage_a()
                stage_b(*out[1:])
                return out[0]
            return pretraining_step(model, sx, sy, sw)

        import os
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
        if self.split:
            with torch.cuda.graph(self.graph, **cap):
                out = stage_a()
            self.loss = out[0]
            self.graph_b = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph_b, pool=self.graph.pool()):
                stage_b(*out[1:])
            del out
            self.bucket_a = FlatGradBucket(upper)
            self.bucket_b = FlatGradBucket(lower)
        else:
            with torch.cuda.graph(self.graph, **cap):
                self.loss = pretraining_step(model, sx, sy, sw)
            if distributed:        # the captured backward fixed the gradient addresses: build the flat bucket over them
                self.bucket = FlatGradBucket(model.parameters())

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

    def __call__(self, X, Y, W, prefetch=None):
        """One step on the batch (X, Y, W).  ``prefetch=(Xn, Yn, Wn)`` starts copying the NEXT batch while this one runs."""
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
        self.graph.replay()
        if prefetch is not None:
            self.prefetch(*prefetch)
        if self.split:
            self.bucket_a.gather()
            work = self.bucket_a.all_reduce(async_op=True)       # NCCL on its own stream ...
            self.graph_b.replay()                                # ... while the second half of backward runs
            self.bucket_b.gather()
            self.bucket_b.all_reduce()
            if work is not None:
                work.wait()
            self.bucket_a.scatter()
            self.bucket_b.scatter()
        elif self.bucket is not None:
            self.bucket.gather()
            self.bucket.all_reduce()
            self.bucket.scatter()
        return self.loss


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


def _lib_error(msg):
    from ._lib import PBError
    return PBError(msg)
