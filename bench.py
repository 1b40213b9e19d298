#!/usr/bin/env python
"""bench.py -- ProteinBERT pretraining throughput (fwd + loss + bwd) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One JSON line on stdout (rank 0).  A "step" is one forward + loss + backward over one synthetic batch of
BASELINE.json configs[1] (batch 64 per GPU, seq 512, 8943 GO labels, 6 blocks, local 128 / global 512, bf16),
plus the gradient all-reduce when N > 1 (weak scaling: the per-GPU batch is fixed).

  value        samples/s over all ranks, inputs resident in HBM, CUDA-event timed, max over ranks
  e2e          same metric through the public API from pinned HOST buffers (H2D of the six input tensors and
               a D2H read of the loss inside the timed region)
  roofline     the dominant kernel (conv pair forward): algorithmic FLOPs per launch / its mean launch time,
               measured live with CUDA events, against MEASURED_PEAKS.json
  cpu_baseline the CPU oracle port of the reference on this box's host cores, bounded sample (rank 0, N=1)
--impl reference times that CPU port alone (the reference is pure Python/torch and cannot travel to the box).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(sequences_length=512, num_annotations=8943, local_dim=128, global_dim=512, key_dim=64, num_heads=4, num_blocks=6)
PER_GPU_BATCH = 64
METRIC = "pretrain samples/sec (seq512, fwd+bwd)"
WORKLOAD = "configs[1]: ProteinBERT pretraining step, batch 64 per GPU, seq 512, 8943 GO labels, 6 blocks 128/512"
CPU_SAMPLE_BATCH = 8          # configs[0]: the reference's own CPU-runnable case (batch 8, seq 512)


def fallback_peaks():
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "_source": "fallback"}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        d["_source"] = "measured"
        return d
    return fallback_peaks()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.proc, self.lines, self.idx = None, [], gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": (sm[len(sm) // 2] if sm else None), "sm_max_mhz": (max(mx) if mx else None),
                "samples": len(sm), "reasons": sorted(reasons)}


def flops_per_sample():
    """algorithmic fwd+bwd FLOPs per sample as the reference executes it."""
    L, C, G, A, V, K, H, NB = 512, 128, 512, 8943, 26, 64, 4, 6
    blk = 2 * (2 * L * C * C * 9) + 2 * G * C + 2 * L * C * C + 2 * L * C * G + 2 * L * C * K * H + 2 * K * K * G * H \
        + 2 * K * K * L * H + 2 * K * L * G + 2 * K * G + 2 * 2 * G * G
    fwd = NB * blk + 2 * A * G + 2 * L * C * V + 2 * G * A
    return fwd + 2 * (fwd - 2 * A * G) + 2 * A * G


def cpu_port_baseline(max_seconds=20.0, min_steps=3, max_steps=40, batch=CPU_SAMPLE_BATCH):
    """The CPU restatement of the reference (oracle/, torch CPU fp32, all host threads): fwd+loss+bwd on a
    bounded sample of the workload -- batch 8 of the same seq-512 / 8943-label / 6-block model."""
    import torch
    from oracle import model_oracle
    torch.set_num_threads(os.cpu_count() or 1)
    params = model_oracle.init_params(seed=0, **CFG)
    b = model_oracle.synthetic_batch(batch, CFG["sequences_length"], CFG["num_annotations"], seed=0)
    import warnings
    warnings.simplefilter("ignore")
    for _ in range(2):
        model_oracle.loss_and_grads(params, b, attention="as_executed")
    times = []
    t_all = time.perf_counter()
    while len(times) < max_steps and (len(times) < min_steps or time.perf_counter() - t_all < max_seconds):
        t0 = time.perf_counter()
        model_oracle.loss_and_grads(params, b, attention="as_executed")
        times.append(time.perf_counter() - t0)
    times.sort()
    med = times[len(times) // 2]
    return {"value": batch / med, "unit": "samples/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"oracle/model_oracle.py (torch CPU fp32, attention as executed), batch {batch} x seq 512, "
                      f"fwd+loss+bwd, median of {len(times)} steps after 2 warm-up", "ms_per_step": med * 1e3}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    base = cpu_port_baseline(max_seconds=max(5.0, 2.0 * args.steps), min_steps=min(args.steps, 5), max_steps=max(args.steps, 5))
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": base["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "global_batch": CPU_SAMPLE_BATCH, "seq_len": 512,
                       "note": "the reference is pure Python/torch and is not present on the GPU box; this arm times its "
                               "CPU restatement (oracle/) on the host cores, batch 8 per step"},
            "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": base["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--graph", type=int, default=1, help="capture the step in a CUDA graph for the `value` leg (N=1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist
    import proteinbert_b200 as pb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    pb._lib.check(pb._lib.lib().pb_device_ok(local_rank), "pb_device_ok")
    if world > 1:
        _opts = dist.ProcessGroupNCCL.Options()
        _opts.is_high_priority_stream = True      # the all-reduce has to get SMs while backward kernels are queued
        dist.init_process_group("nccl", device_id=dev, pg_options=_opts)

    torch.manual_seed(0)
    model = pb.ProteinBERT(device=dev, precision=args.precision, **CFG)
    reducer = None
    if world > 1:
        reducer = pb.GradAllReducer(model.parameters())
        reducer.broadcast_parameters(0)
        for t in model.head_tensors().values():        # unregistered in the reference: broadcast once by hand
            dist.broadcast(t.data, 0)
    B = PER_GPU_BATCH
    host = pb.data_processing.synthetic_batch(B, CFG["sequences_length"], CFG["num_annotations"], seed=100 + rank)
    host = [t.pin_memory() for t in host]
    devb = [t.to(dev) for t in host]
    h2d_bytes = sum(t.numel() * t.element_size() for t in host)

    def dicts(b):
        return ({"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]})

    X, Y, W = dicts(devb)

    def step():
        return pb.pretraining_step(model, X, Y, W, reducer)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (also sets kernel attributes and sizes the split-K scratch) ---------------------------------
    for _ in range(args.warmup):
        step()
    barrier()

    # ---- CUDA graph of forward+loss+backward, the package's public fast path; for N > 1 every call adds the
    #      flat-bucket gradient all-reduce (gather kernel, NCCL all-reduce, scatter kernel) after the replay ----------
    graph = None
    launches_per_step = None
    if args.graph:
        try:
            n0 = pb.launch_count()
            graph = pb.GraphedPretrainStep(model, X, Y, W, warmup=1)
            launches_per_step = (pb.launch_count() - n0) // 2          # one warm-up step + the captured step
            graph(X, Y, W)
            torch.cuda.synchronize()
        except Exception as e:          # capture is an optimisation of launch overhead, not a correctness path
            print(f"[bench] CUDA graph capture failed ({type(e).__name__}: {e}); timing eagerly", file=sys.stderr)
            graph = None
            torch.cuda.synchronize()

    # ---- value: K steps, device-resident inputs ---------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    n0 = pb.launch_count()
    e0.record()
    for _ in range(args.steps):
        if graph is not None and world == 1:
            graph.graph.replay()
        elif graph is not None:
            graph(X, Y, W)               # replay + gradient bucket all-reduce
        else:
            step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = ((launches_per_step + (2 if world > 1 else 0)) * args.steps) if graph is not None else (pb.launch_count() - n0)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    clocks = sampler.stop() if rank == 0 else None

    # ---- dominant kernel: conv pair forward, timed live with CUDA events around its launches ------------------
    from proteinbert_b200 import ops
    prof = []
    ops.PROFILE = {"pb_conv_pair_fwd": prof}
    for _ in range(max(3, min(args.steps, 10))):
        step()
    torch.cuda.synchronize()
    ops.PROFILE = None
    conv_ms = sorted(a.elapsed_time(b) for a, b in prof)
    conv_ms_eager = sum(conv_ms) / len(conv_ms)      # includes the host time between the event and the launch (idle GPU)
    L, C = CFG["sequences_length"], CFG["local_dim"]
    # the same entry point on config-2 tensors with block 0's weights, launched back to back so that the stream never
    # drains: the event pair then brackets kernel time only
    blk0 = model.proteinBERT_blocks[0]
    adt = torch.bfloat16 if args.precision == "bf16" else torch.float32
    dtc = 1 if args.precision == "bf16" else 0
    xin = torch.randn(B, L, C, device=dev).to(adt)
    packs = [ops.conv_pack(m[0].weight.detach(), dtc) for m in (blk0.local_narrow_conv_layer, blk0.local_wide_conv_layer)]
    bn_, bw_ = blk0.local_narrow_conv_layer[0].bias.detach(), blk0.local_wide_conv_layer[0].bias.detach()
    g2l_ = torch.randn(B, C, device=dev)
    zn_, zw_, u_ = torch.empty_like(xin), torch.empty_like(xin), torch.empty_like(xin)
    st_ = torch.zeros(B, 2, dtype=torch.float64, device=dev)
    cs = torch.cuda.current_stream().cuda_stream

    def conv_once():
        ops._call("pb_conv_pair_fwd", xin.data_ptr(), packs[0][0].data_ptr(), packs[1][0].data_ptr(), bn_.data_ptr(), bw_.data_ptr(),
                  g2l_.data_ptr(), zn_.data_ptr(), zw_.data_ptr(), u_.data_ptr(), st_.data_ptr(), B, L, C, blk0.conv_kernel_size,
                  blk0.wide_conv_dilation, dtc, cs)

    for _ in range(5):
        conv_once()
    n_conv = 60
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    c0.record()
    for _ in range(n_conv):
        conv_once()
    c1.record()
    torch.cuda.synchronize()
    conv_ms_mean = c0.elapsed_time(c1) / n_conv
    conv_flops = 2 * (2.0 * B * L * C * C * 9)            # two convolutions, one launch (19.33 GF at B=64)
    peaks = load_peaks()
    peak_tf = peaks["bf16_tflops_sustained"]
    achieved_tf = conv_flops / (conv_ms_mean * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r01_conv_traffic.json")
    if os.path.exists(tpath) and B == 64:          # dram bytes per launch from the committed ncu --set full capture
        t = json.load(open(tpath))
        traffic = t["dram_bytes_read"] + t["dram_bytes_write"]
    roofline = {"kernel": "conv_pair_fwd", "bound": "tensor", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": achieved_tf / peak_tf, "traffic": traffic, "peak_source": f"{peaks['_source']} bf16_tflops_sustained",
                "us_per_launch": conv_ms_mean * 1e3, "launches_timed": n_conv,
                "timing": "CUDA events around 60 back-to-back launches of pb_conv_pair_fwd on config-2 tensors (block 0 weights)",
                "us_per_launch_in_eager_step": conv_ms_eager * 1e3}

    # ---- e2e: public API from pinned host buffers, H2D of inputs + D2H of the loss inside the timed region ----
    def e2e_step():
        if graph is not None:            # public fast path: H2D into the static buffers, graph replay, D2H of the loss
            Xh, Yh, Wh = dicts(host)
            # the next batch's H2D (here: the same pinned host tensors again, a full 9.9 MB copy per step) is started on a
            # side stream while this step computes -- what a data loader with one batch of look-ahead does
            return graph(Xh, Yh, Wh, prefetch=(Xh, Yh, Wh)).item()
        b = [t.to(dev, non_blocking=True) for t in host]
        Xh, Yh, Wh = dicts(b)
        loss = pb.pretraining_step(model, Xh, Yh, Wh, reducer)
        return loss.item()

    for _ in range(3):
        e2e_step()
    barrier()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t[0])

    # ---- optimizer step (a20 / f-2), reported separately: it is outside the metric's timed region -----------
    adam = None
    if world == 1:
        opt = pb.FusedAdam(model.parameters(), lr=2e-4)
        for _ in range(3):
            opt.step()
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(20):
            opt.step()
        a1.record()
        torch.cuda.synchronize()
        n_par = sum(p.numel() for p in model.parameters() if p.grad is not None)
        us_api = a0.elapsed_time(a1) / 20 * 1e3          # FusedAdam.step() back to back: bounded by its Python host code
        tab = opt._tables[0]                              # the kernel alone: raw launches on the optimizer's own table
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k0.record()
        for i in range(20):
            ops._call("pb_adam_step", tab["table"].data_ptr(), tab["rows"], 2e-4, 0.9, 0.999, 1e-8, 0.0, tab["t"] + 1 + i,
                      torch.cuda.current_stream().cuda_stream)
        k1.record()
        torch.cuda.synchronize()
        us = k0.elapsed_time(k1) / 20 * 1e3
        adam = {"kernel": "pb_adam_step (one launch for all 133 tensors)", "us_per_step": us, "us_per_FusedAdam_step_call": us_api,
                "bytes": 28 * n_par,
                "GBps": 28 * n_par / (us * 1e-6) / 1e9, "frac_of_hbm_peak": 28 * n_par / (us * 1e-6) / 1e9 / load_peaks()["hbm_gbs"],
                "note": "outside the timed region of `value` and `e2e`"}

    if rank == 0:
        gb = B * world
        value = gb * args.steps / (ms * 1e-3)
        line = {"metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "global_batch": gb, "seq_len": CFG["sequences_length"],
                           "parallelism": f"dp{world}", "cuda_graph": graph is not None,
                           "l2": "no explicit flush: one step touches ~1 GB of activations and weights, 8x the 126 MB L2",
                           "timed_region": "forward + loss + backward" + (" + gradient all-reduce" if world > 1 else "")},
                "clocks": clocks,
                "e2e": {"value": gb * args.steps / (ms_e2e * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": h2d_bytes,
                        "d2h_bytes_per_step": 8, "ms_per_step": ms_e2e / args.steps},
                "gpu_launches": int(launches), "roofline": roofline,
                "model_tflops": flops_per_sample() * value / 1e12}
        if adam is not None:
            line["adam_step"] = adam
        if world == 1 and not args.no_cpu_baseline:
            base = cpu_port_baseline()
            line["cpu_baseline"] = {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
