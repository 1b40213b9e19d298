#!/usr/bin/env python
"""bench.py -- ProteinBERT pretraining throughput (fwd + loss + bwd) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config 2|4|5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One JSON line on stdout (rank 0).  A "step" is one forward + loss + backward over one synthetic batch of the chosen
BASELINE.json configuration (default configs[1]: batch 64 per GPU, seq 512, 8943 GO labels, 6 blocks, local 128 /
global 512, bf16), plus the gradient all-reduce when N > 1 (weak scaling: the per-GPU batch is fixed).

  value               samples/s over all ranks, inputs resident in HBM, CUDA-event timed, max over ranks
  e2e                 same metric through the public API from pinned HOST buffers (H2D of the six input tensors and
                      a D2H read of the loss inside the timed region)
  roofline            the dominant kernel (conv pair forward): algorithmic FLOPs per launch / its mean launch time,
                      measured live with CUDA events, against MEASURED_PEAKS.json
  roofline_table      every kernel of one step: launches, mean duration (CUPTI activity records taken live in this
                      process over eager steps after the timed region), algorithmic FLOPs / bytes, floor, fraction;
                      plus the corruption kernel and the fused Adam step timed with CUDA events
  cpu_baseline        the reference's own modules.py + the loss of utils.py:293-294 on this box's host cores (from
                      baseline/_ref, staged by scripts/stage_reference.sh; the oracle port if absent), bounded sample
  gpu_eager_baseline  the same unmodified reference modules on cuda:0 at the workload's batch: fp32 with TF32 off,
                      TF32 on, bf16 autocast (what eager torch reaches through cuDNN / cuBLAS), and the cuDNN conv pair
--impl reference times the CPU reference alone on the same model configuration.
"""
import argparse
import importlib.util
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time
import warnings

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BASE_CFG = dict(sequences_length=512, num_annotations=8943, local_dim=128, global_dim=512, key_dim=64, num_heads=4, num_blocks=6)
CONFIGS = {
    2: dict(cfg=dict(BASE_CFG), per_gpu_batch=64, metric="pretrain samples/sec (seq512, fwd+bwd)",
            workload="configs[1]: ProteinBERT pretraining step, batch 64 per GPU, seq 512, 8943 GO labels, 6 blocks 128/512"),
    4: dict(cfg=dict(BASE_CFG, sequences_length=2048), per_gpu_batch=16, metric="pretrain samples/sec (seq2048, fwd+bwd)",
            workload="configs[3]: long-sequence pretraining step, batch 16 per GPU, seq 2048, 8943 GO labels, 6 blocks 128/512"),
    5: dict(cfg=dict(BASE_CFG, sequences_length=1024, local_dim=256, global_dim=1024, num_blocks=12), per_gpu_batch=32,
            metric="pretrain samples/sec (seq1024, fwd+bwd)",
            workload="configs[4]: scaled-width pretraining step, local 256 / global 1024, 12 blocks, seq 1024, batch 32 per GPU "
                     "(assumed: key 64, 4 heads, 8943 GO labels -- BASELINE.json leaves them open)"),
}
CPU_SAMPLE_BATCH = 8          # configs[0]: the reference's own CPU-runnable case (batch 8)
REF_DIR = os.path.join(ROOT, "baseline", "_ref")


def fallback_peaks():
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "_source": "fallback"}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        d["_source"] = "measured"
        d.setdefault("bf16_tflops_sustained", d.get("bf16_tflops"))
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


def flops_per_sample(cfg, as_executed_by_reference=False):
    """algorithmic fwd+bwd FLOPs per sample.  The kernels evaluate the attention layer in its
    closed form (the reference's Softmax(dim=1) makes every weight 1/K), which drops the Q/K/score/PV products; the
    larger figure counts those as the reference executes them."""
    L, C, G, A = cfg["sequences_length"], cfg["local_dim"], cfg["global_dim"], cfg["num_annotations"]
    K, H, NB, V = cfg["key_dim"], cfg["num_heads"], cfg["num_blocks"], 26
    blk = 2 * (2 * L * C * C * 9) + 2 * G * C + 2 * L * C * C + 2 * L * C * G + 2 * 2 * G * G
    if as_executed_by_reference:
        blk += 2 * L * C * K * H + 2 * K * K * G * H + 2 * K * K * L * H + 2 * K * L * G + 2 * K * G
    fwd = NB * blk + 2 * A * G + 2 * L * C * V + 2 * G * A
    return fwd + 2 * (fwd - 2 * A * G) + 2 * A * G


# ------------------------------------------------------------------------------------------------------------------
# the reference itself (baseline/_ref: unmodified copies staged by scripts/stage_reference.sh), CPU and eager GPU
# ------------------------------------------------------------------------------------------------------------------
def load_reference_modules():
    path = os.path.join(REF_DIR, "modules.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("pb_reference_modules", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def reference_step_fn(ref, cfg, batch, device, autocast_dtype=None):
    """One iteration of the reference loop body, utils.py:291-299 (model(X), the loss expression, zero_grad, backward),
    on the unmodified reference ProteinBERT; the loss modules are those of dummy_tests.py:132-133."""
    import torch
    torch.manual_seed(0)
    model = ref.ProteinBERT(device=device, **cfg)
    model.train()
    ce, bce = torch.nn.CrossEntropyLoss(reduction="none"), torch.nn.BCELoss(reduction="none")
    xl, xg, yl, yg, wl, wg = [t.to(device) for t in batch]
    X = {"local": xl, "global": xg}

    def step():
        for p in model.parameters():
            p.grad = None
        if autocast_dtype is not None:
            with torch.autocast("cuda", dtype=autocast_dtype):
                pl, pg = model(X)
            pl, pg = pl.float(), pg.float()         # BCELoss refuses to run under autocast
        else:
            pl, pg = model(X)
        loss = torch.mean(ce(pl.permute(0, 2, 1), yl) * wl) + torch.mean(bce(pg, yg.float()) * wg)
        loss.backward()
        return loss

    return step


def cpu_reference_baseline(cfg, max_seconds=20.0, min_steps=3, max_steps=40, batch=CPU_SAMPLE_BATCH):
    """fwd + loss + bwd of the reference on the host cores, bounded sample (batch 8 of the same model)."""
    import torch
    import proteinbert_b200 as pb
    torch.set_num_threads(os.cpu_count() or 1)
    warnings.simplefilter("ignore")
    L, A = cfg["sequences_length"], cfg["num_annotations"]
    ref = load_reference_modules()
    if ref is not None:
        b = pb.data_processing.synthetic_batch(batch, L, A, seed=0)
        step = reference_step_fn(ref, cfg, b, "cpu")
        kind = "reference"
        what = "baseline/_ref/modules.py (unmodified reference, torch CPU fp32) + loss of utils.py:293-294"
    else:
        from oracle import model_oracle
        params = model_oracle.init_params(seed=0, **cfg)
        b = model_oracle.synthetic_batch(batch, L, A, seed=0)
        step = lambda: model_oracle.loss_and_grads(params, b, attention="as_executed")
        kind = "port"
        what = "oracle/model_oracle.py (CPU restatement; baseline/_ref not staged), torch CPU fp32, attention as executed"
    for _ in range(2):
        step()
    times = []
    t_all = time.perf_counter()
    while len(times) < max_steps and (len(times) < min_steps or time.perf_counter() - t_all < max_seconds):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
    times.sort()
    med = times[len(times) // 2]
    cpu_model = ""
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                cpu_model = ln.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    return {"value": batch / med, "unit": "samples/s", "cores": torch.get_num_threads(), "kind": kind, "cpu": cpu_model,
            "sample": f"{what}; batch {batch} x seq {L}, fwd+loss+bwd, median of {len(times)} steps after 2 warm-up",
            "ms_per_step": med * 1e3}


def gpu_eager_baseline(cfg, B, dev, steps=8):
    """The 'library kernels to beat' rows: the unmodified reference modules on this GPU through eager
    torch (cuDNN convolutions, cuBLASLt GEMMs, ATen element-wise kernels), and the cuDNN convolution pair alone."""
    import torch
    import proteinbert_b200 as pb
    ref = load_reference_modules()
    out = {"batch": B, "timing": f"CUDA events around {steps} steps after 3 warm-up, cudnn.benchmark on"}
    if ref is None:
        out["unavailable"] = "baseline/_ref not staged (scripts/stage_reference.sh)"
        return out
    warnings.simplefilter("ignore")
    L, A, C = cfg["sequences_length"], cfg["num_annotations"], cfg["local_dim"]
    batch = pb.data_processing.synthetic_batch(B, L, A, seed=0)
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32, torch.backends.cudnn.benchmark)
    torch.backends.cudnn.benchmark = True

    def timed(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / n

    try:
        for name, tf32, ac in (("fp32_tf32_off", False, None), ("fp32_tf32_on", True, None), ("bf16_autocast", True, torch.bfloat16)):
            torch.backends.cuda.matmul.allow_tf32 = tf32
            torch.backends.cudnn.allow_tf32 = tf32
            try:
                step = reference_step_fn(ref, cfg, batch, dev, ac)
                ms = timed(step, steps)
                out[name] = {"ms_per_step": ms, "samples_per_s": B / (ms * 1e-3)}
                del step
            except Exception as e:          # a baseline row, never a reason to lose the bench line
                out[name] = {"error": f"{type(e).__name__}: {e}"[:200]}
            torch.cuda.empty_cache()
        # the convolution pair alone: nn.Conv1d k=9 d=1 and d=5, 'same', channels-first as the reference calls them
        conv = {}
        for name, dt, tf32 in (("bf16", torch.bfloat16, True), ("tf32", torch.float32, True), ("fp32", torch.float32, False)):
            torch.backends.cudnn.allow_tf32 = tf32
            x = torch.randn(B, C, L, device=dev, dtype=dt, requires_grad=True)
            cn = torch.nn.Conv1d(C, C, 9, padding="same", device=dev, dtype=dt)
            cw = torch.nn.Conv1d(C, C, 9, padding="same", dilation=5, device=dev, dtype=dt)
            with torch.no_grad():
                fwd_ms = timed(lambda: (cn(x), cw(x)), 30)
            go = torch.randn(B, C, L, device=dev, dtype=dt)

            def fb():
                yn, yw = cn(x), cw(x)
                torch.autograd.grad([yn, yw], [x, cn.weight, cw.weight], [go, go])

            fb_ms = timed(fb, 20)
            conv[name] = {"fwd_us": fwd_ms * 1e3, "fwd_bwd_us": fb_ms * 1e3,
                          "fwd_tflops": 2 * (2.0 * B * L * C * C * 9) / (fwd_ms * 1e-3) / 1e12}
        out["cudnn_conv_pair"] = conv
        out["cudnn_conv_pair"]["note"] = ("two torch.nn.Conv1d calls (no bias-GELU-add-LayerNorm-sum epilogue, which pb_conv_pair_fwd "
                                          "includes); fwd_bwd adds dgrad + wgrad through autograd")
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32, torch.backends.cudnn.benchmark = old
        torch.cuda.empty_cache()
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    conf = CONFIGS[args.config]
    base = cpu_reference_baseline(conf["cfg"], max_seconds=max(5.0, 2.0 * args.steps), min_steps=min(args.steps, 5),
                                  max_steps=max(args.steps, 5))
    line = {"impl": "reference", "metric": conf["metric"], "value": base["value"], "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": base["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": conf["workload"], "global_batch": CPU_SAMPLE_BATCH, "seq_len": conf["cfg"]["sequences_length"],
                       "note": "each step is a bounded sample of the workload: batch 8 of the same model on the host cores "
                               "(configs[0] is exactly this case for the default workload)"},
            "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "cpu")},
            "e2e": {"value": base["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# per-kernel roofline table
# ------------------------------------------------------------------------------------------------------------------
def kernel_specs(cfg, B):
    """(label, regex on the demangled kernel name, grid predicate or None, bound, algorithmic GFLOP, algorithmic MB).
    Algorithmic = compulsory work of the op at this configuration: each tensor
    read or written once in its storage type.  act = one [B, L, C] bf16 activation."""
    L, C, G, A, V = cfg["sequences_length"], cfg["local_dim"], cfg["global_dim"], cfg["num_annotations"], 26
    act = B * L * C * 2 / 1e6
    conv_gf = 2 * (2.0 * B * L * C * C * 9) / 1e9
    pool_gf = 2.0 * B * L * C * G / 1e9
    lin_gf = 2.0 * B * L * C * C / 1e9
    dzv = B * L * G * 2 / 1e6
    wA = A * G * 4 / 1e6
    gw = (G * G + G * C) * 4 / 1e6
    return [
        ("conv pair fwd (+bias, GELU, 4-way add, LN sums)", r"tc_conv2?_kernel<(\(bool\))?(1|true)", None, "tensor", conv_gf, 4 * act),
        ("conv pair data gradient", r"tc_conv2?_kernel<(\(bool\))?(0|false)", None, "tensor", conv_gf, 4 * act),
        ("conv pair weight gradient", r"tc_conv_wgrad3_kernel", None, "tensor", conv_gf, 3 * act),
        ("conv weight-gradient split-K reduce", r"conv_wgrad_reduce_kernel", None, "hbm", 0.0, 2 * 9 * C * C * 4 / 1e6),
        ("attention pooling fwd: sum_l gelu(h Wv)", r"tc_pool_kernel<(\(int\))?0", None, "tensor", pool_gf, act),
        ("attention pooling bwd: dzv = dpool gelu'(h Wv)", r"tc_pool_kernel<(\(int\))?1", None, "hbm", pool_gf, act + dzv),
        ("attention dh = dh_in + dzv Wv^T (K = G)", r"tc_gemm_kernel<(\(int\))?3", None, "hbm", pool_gf, dzv + 2 * act),
        ("local linear dh = du2 + dz W (K = C)", r"tc_gemm_kernel<(\(int\))?3", None, "hbm", lin_gf, 3 * act),
        ("local linear fwd (+bias, GELU, residual, LN sums)", r"tc_gemm_kernel<(\(int\))?0", None, "hbm", lin_gf, 3 * act),
        ("weight gradients dW2 / dWv / dWo (MN-major GEMM)", r"tc_wgrad_kernel", None, "tensor", (lin_gf + pool_gf) / 2, act + dzv / 2),
        ("split-K reduce of those", r"(?<!conv_)wgrad_reduce_kernel", None, "hbm", 0.0, (C * C + G * C) * 4 / 2e6),
        ("joint LayerNorm apply", r"ln_apply_bf16_kernel|ln_apply_kernel", None, "hbm", 0.0, 2 * act + 2 * L * C * 4 / 1e6),
        ("joint LayerNorm bwd: per-sample sums", r"ln_bwd_sums_vec_kernel", None, "hbm", 0.0, 2 * act),
        ("joint LayerNorm bwd + two GELU bwd (LN1)", r"ln_bwd_apply_gelu_kernel<.*, ?(\(bool\))?(1|true)>", None, "hbm", 0.0, 7 * act),
        ("joint LayerNorm bwd + one GELU bwd (LN2)", r"ln_bwd_apply_gelu_kernel<.*, ?(\(bool\))?(0|false)>", None, "hbm", 0.0, 5 * act),
        ("global track", r"global_\w+_kernel|gtrack_\w+", None, "hbm", 0.0, gw + 6 * B * G * 4 / 1e6),
        ("global track weight gradients dW = dZ^T X (side stream)", r"gemm_simt_kernel<.*ColMajorA<float>.*EpStore", None, "hbm", 2.0 * B * G * G / 1e9, (G * G + 2 * B * G) * 4 / 1e6),
        ("annotation layers (GO in / GO out, fwd + dX + dW)", r"tc_skinny_kernel", None, "hbm", 2.0 * B * A * G / 1e9, wA),
        ("annotation split-K reduce / dz", r"skinny_reduce_kernel|skinny_dz_kernel", None, "hbm", 0.0, wA / 4),
        ("token head logits / dh / dWo", r"tc_gemm_kernel<(\(int\))?[45]", None, "hbm", 2.0 * B * L * C * 128 / 1e9, act + B * L * 32 * 4 / 1e6),
        ("token head batch-axis softmax / dlogit", r"softmax_batch_kernel|token_dlogit_kernel", None, "hbm", 0.0, 2 * B * L * V * 4 / 1e6),
        ("loss (double-softmax CE + clamped BCE, fp64)", r"loss_kernel", None, "hbm", 0.0, (2 * B * L * V * 4 + 2 * B * A * 4 + B * A * 12 + B * L * 16) / 1e6),
        ("embedding gather / scatter-add", r"embedding_", None, "hbm", 0.0, act + B * L * 8 / 1e6),
        ("weight staging (fp32 masters -> bf16 packs)", r"stage_weights_kernel", None, "hbm", 0.0,
         cfg["num_blocks"] * (2 * 9 * C * C + C * C + C * G) * (4 + 2 * 2) / 1e6),
    ]


def collect_kernel_table(step_fn, n_steps, cfg, B, peaks):
    """Run eager steps under torch.profiler (CUPTI activity records: hardware start/end of every kernel, graph or not)
    and fold them by kernel into launches/step, mean microseconds, algorithmic work, floor and fraction of the floor."""
    import torch
    from torch.profiler import ProfilerActivity, profile
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(n_steps):
            step_fn()
        torch.cuda.synchronize()
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "trace.json")
        prof.export_chrome_trace(path)
        trace = json.load(open(path))
    kernels = [e for e in trace.get("traceEvents", []) if e.get("cat") == "kernel" and "dur" in e]
    specs = kernel_specs(cfg, B)
    rows = {}
    for e in kernels:
        name = e["name"]
        grid = e.get("args", {}).get("grid", [0, 0, 0])
        label, spec = None, None
        for s in specs:
            if re.search(s[1], name):
                if s[0].startswith("attention dh"):          # same kernel as the local linear's dh; told apart by K
                    continue
                label, spec = s[0], s
                break
        key = label or name[:90]
        r = rows.setdefault(key, {"kernel": key, "name": name[:120], "launches": 0, "us_total": 0.0, "spec": spec, "grids": set()})
        r["launches"] += 1
        r["us_total"] += float(e["dur"])
        r["grids"].add(tuple(grid))
    # the ADD_STORE GEMM serves two ops (local linear dh, K = C; attention dh, K = G) with the same grid, which the trace
    # cannot tell apart: they share one row whose algorithmic work is the mean of one launch of each
    tf_peak, bw_peak = peaks["bf16_tflops_sustained"], peaks["hbm_gbs"]
    table, total_us, total_floor = [], 0.0, 0.0
    for key, r in rows.items():
        per_step = r["launches"] / n_steps
        us = r["us_total"] / r["launches"]
        row = {"kernel": key, "launches_per_step": round(per_step, 2), "us_per_launch": round(us, 2),
               "us_per_step": round(r["us_total"] / n_steps, 1)}
        spec = r["spec"]
        if spec is not None:
            gf, mb = spec[4], spec[5]
            if spec[0].startswith("local linear dh"):        # half of these launches are the attention dh GEMM (K = G)
                other = [s for s in specs if s[0].startswith("attention dh")][0]
                gf, mb = (gf + other[4]) / 2, (mb + other[5]) / 2
                row["kernel"] = "dh GEMMs: local linear (K = C) and attention (K = G), mean of both"
            floor_tc = gf * 1e9 / (tf_peak * 1e12) * 1e6
            floor_bw = mb * 1e6 / (bw_peak * 1e9) * 1e6
            floor = max(floor_tc, floor_bw)
            row.update({"bound": "tensor" if floor_tc >= floor_bw else "hbm", "alg_gflop": round(gf, 3), "alg_mb": round(mb, 2),
                        "floor_us": round(floor, 2), "frac": round(floor / us, 3) if us > 0 else None})
            total_floor += floor * per_step
        else:
            row.update({"bound": None, "frac": None, "name": r["name"]})
        total_us += r["us_total"] / n_steps
        table.append(row)
    table.sort(key=lambda r: -r["us_per_step"])
    return {"rows": table, "kernel_us_per_step_serialised": round(total_us, 1), "floor_us_per_step": round(total_floor, 1),
            "frac_of_floors": round(total_floor / total_us, 3) if total_us else None,
            "launches_per_step": round(sum(r["launches"] for r in rows.values()) / n_steps, 1),
            "peaks": {"bf16_tflops_sustained": tf_peak, "hbm_gbs": bw_peak, "source": peaks["_source"]},
            "timing": f"CUPTI kernel activity records (torch.profiler) over {n_steps} eager steps in this process, after the timed "
                      "region; in-step durations, so fractions are against the SUSTAINED peaks; side-stream overlap is not "
                      "reflected in the serialised sum"}


def time_corrupt_pad(pb, B, L, A, dev, peaks):
    """pb_corrupt_pad (a19) on a synthetic ragged batch of the workload's shape: CUDA events around back-to-back launches."""
    import random
    import numpy as np
    import torch
    rnd = random.Random(7777)
    letters = pb.data_processing.ALL_AMINO_ACIDS
    seqs = ["".join(letters[rnd.randint(0, 21)] for _ in range(rnd.randint(0, 2 * L))) for _ in range(B)]
    anns = (np.random.RandomState(0).rand(B, A) < 0.005).astype(np.int64)
    torch.manual_seed(0)
    draws = [pb.data_processing.draw_sample(len(s), A, L) for s in seqs]
    packed = pb.data_processing.pack_batch(seqs, anns, draws, L)
    d = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in packed.items()}
    out = None
    for _ in range(3):
        out = pb.data_processing.corrupt_pad(d, dev, out=out)
    torch.cuda.synchronize()
    n = 50
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        pb.data_processing.corrupt_pad(d, dev, out=out)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) / n * 1e3
    in_bytes = sum(v.numel() * v.element_size() for v in d.values() if isinstance(v, torch.Tensor))
    out_bytes = B * L * (8 + 8 + 8) + B * A * (4 + 4 + 8)
    mb = (in_bytes + out_bytes) / 1e6
    floor = mb * 1e6 / (peaks["hbm_gbs"] * 1e9) * 1e6
    return {"kernel": "pb_corrupt_pad (a19; outside the step)", "launches_per_step": 0, "us_per_launch": round(us, 2), "bound": "hbm",
            "alg_mb": round(mb, 2), "floor_us": round(floor, 2), "frac": round(floor / us, 3),
            "timing": f"CUDA events around {n} back-to-back calls of data_processing.corrupt_pad(out=previous batch) (host launch latency of one small kernel per call included); "
                      "against the burst HBM peak"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS), help="BASELINE.json configuration (2 = configs[1], the metric's)")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch override (e.g. configs[2]: global batch 512 -> 512 / N)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--graph", type=int, default=1, help="capture the step in a CUDA graph for the `value` leg")
    ap.add_argument("--ddp-overlap", type=int, default=1, help="N > 1: two-graph backward with the first all-reduce under the second graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true")
    ap.add_argument("--no-kernel-table", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    args.warmup = max(args.warmup, 3)
    conf = CONFIGS[args.config]
    CFG = conf["cfg"]

    import torch
    import torch.distributed as dist
    import proteinbert_b200 as pb
    from proteinbert_b200 import ops

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
        reducer.broadcast_parameters(0, extra_tensors=model.head_tensors().values())
    B = args.batch or conf["per_gpu_batch"]
    L, C, A = CFG["sequences_length"], CFG["local_dim"], CFG["num_annotations"]
    host = pb.data_processing.synthetic_batch(B, L, A, seed=100 + rank)
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
    #      gradient all-reduce -------------------------------------------------------------------------------------
    graph = None
    launches_per_step = None
    if args.graph:
        try:
            n0 = pb.launch_count()
            graph = pb.GraphedPretrainStep(model, X, Y, W, warmup=1, overlap=bool(args.ddp_overlap))
            launches_per_step = (pb.launch_count() - n0) // 2          # one warm-up step + the captured step
            for _ in range(max(1, min(3, args.warmup))):      # warm replays of the very thing that is timed (first exchanges set up links)
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
    import gc
    gc.collect()
    gc.disable()          # a collection pause on one of N host processes would be charged to every rank's step (max over ranks)
    barrier()
    n0 = pb.launch_count()
    e0.record()
    for _ in range(args.steps):
        if graph is not None and world == 1:
            graph.graph.replay()
        elif graph is not None:
            graph(X, Y, W)               # replay + gradient all-reduce
        else:
            step()
    e1.record()
    barrier()
    gc.enable()
    ms = e0.elapsed_time(e1)
    extra = getattr(graph, "launches_outside_graph", 0) if (graph is not None and world > 1) else 0
    launches = ((launches_per_step + extra) * args.steps) if graph is not None else (pb.launch_count() - n0)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    clocks = sampler.stop() if rank == 0 else None

    # ---- data-parallel check: mean over ranks of the rank-local gradients == what the step leaves in .grad --------
    ddp_check = None
    comm = None
    if world > 1 and graph is not None and hasattr(graph, "ddp_check"):
        ddp_check = graph.ddp_check(X, Y, W)
        if hasattr(graph, "comm_profile"):
            comm = graph.comm_profile(X, Y, W, steps=max(5, min(args.steps, 20)))
            comm["exchange"] = graph.bucket.exchange
            comm["ddp_mode"] = graph.mode
            comm["exchanges_per_step"] = 1 + len(graph.cut_offsets or [])
            comm["early_exchange_bytes"] = [4 * o for o in (graph.cut_offsets or [])]

    # ---- dominant kernel: conv pair forward, timed live with CUDA events around its launches ------------------
    peaks = load_peaks()
    adt = torch.bfloat16 if args.precision == "bf16" else torch.float32
    dtc = 1 if args.precision == "bf16" else 0
    blk0 = model.proteinBERT_blocks[0]
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
    conv_flops = 2 * (2.0 * B * L * C * C * 9)            # two convolutions, one launch (19.33 GF at B=64, L=512)
    # the kernel is timed alone, back to back: the burst figure is the denominator; the sustained one is given beside it
    peak_tf = peaks.get("bf16_tflops", peaks["bf16_tflops_sustained"])
    achieved_tf = conv_flops / (conv_ms_mean * 1e-3) / 1e12
    traffic = None
    for tname in ("r02_conv_traffic.json", "r01_conv_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", tname)
        if os.path.exists(tpath) and B == 64 and L == 512:          # dram bytes per launch from the committed ncu --set full capture
            t = json.load(open(tpath))
            traffic = t["dram_bytes_read"] + t["dram_bytes_write"]
            break
    roofline = {"kernel": "conv_pair_fwd", "bound": "tensor", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": achieved_tf / peak_tf, "traffic": traffic, "peak_source": f"{peaks['_source']} bf16_tflops (burst: kernel timed alone)",
                "frac_of_sustained": achieved_tf / peaks["bf16_tflops_sustained"],
                "us_per_launch": conv_ms_mean * 1e3, "launches_timed": n_conv, "algorithmic_gflop_per_launch": conv_flops / 1e9,
                "timing": f"CUDA events around {n_conv} back-to-back launches of pb_conv_pair_fwd on the workload's tensors (block 0 weights)"}

    # ---- e2e: public API from pinned host buffers, H2D of inputs + D2H of the loss inside the timed region ----
    pending = [None]
    losses_read = [0]

    def e2e_step():
        if graph is not None:            # public fast path: H2D into the static buffers, graph replay, D2H of the loss
            Xh, Yh, Wh = dicts(host)
            # the next batch's H2D (here: the same pinned host tensors again, a full copy per step) is started on a
            # side stream while this step computes -- what a data loader with one batch of look-ahead does.  The loss of
            # EVERY step is read on the host (8 bytes, pinned), one step late: the host waits for step k-1's copy after it
            # has issued step k, so the GPU always has the next step queued (a .item() per step leaves it idle meanwhile)
            graph(Xh, Yh, Wh, prefetch=(Xh, Yh, Wh))
            h = graph.read_loss_async()
            if pending[0] is not None:
                pending[0].value()
                losses_read[0] += 1
            pending[0] = h
            return
        b = [t.to(dev, non_blocking=True) for t in host]
        Xh, Yh, Wh = dicts(b)
        loss = pb.pretraining_step(model, Xh, Yh, Wh, reducer)
        loss.item()
        losses_read[0] += 1

    def e2e_drain():
        if pending[0] is not None:
            pending[0].value()
            losses_read[0] += 1
            pending[0] = None

    for _ in range(3):
        e2e_step()
    e2e_drain()
    barrier()
    losses_read[0] = 0
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e2e_drain()                          # the last step's loss is on the host before the region ends
    e1.record()
    barrier()
    assert losses_read[0] == args.steps
    ms_e2e = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t[0])

    # ---- per-kernel table (rank 0, N = 1), the corruption kernel and the optimizer step ---------------------------------
    table = None
    adam = None
    if world == 1:
        if not args.no_kernel_table:
            try:
                table = collect_kernel_table(lambda: pb.pretraining_step(model, X, Y, W), 3, CFG, B, peaks)
                table["rows"].append(time_corrupt_pad(pb, B, L, A, dev, peaks))
            except Exception as e:
                table = {"error": f"{type(e).__name__}: {e}"[:300]}
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
        adam = {"kernel": "pb_adam_step (one launch for all registered tensors)", "us_per_step": us, "us_per_FusedAdam_step_call": us_api,
                "bytes": 28 * n_par,
                "GBps": 28 * n_par / (us * 1e-6) / 1e9, "frac_of_hbm_peak": 28 * n_par / (us * 1e-6) / 1e9 / peaks["hbm_gbs"],
                "note": "outside the timed region of `value` and `e2e`"}
        if table is not None and "rows" in table:
            table["rows"].append({"kernel": "pb_adam_step (a20 / f-2; outside the step)", "launches_per_step": 0, "us_per_launch": round(us, 2),
                                  "bound": "hbm", "alg_mb": round(28 * n_par / 1e6, 1), "floor_us": round(28 * n_par / (peaks["hbm_gbs"] * 1e3), 2),
                                  "frac": round(adam["frac_of_hbm_peak"], 3)})

    if rank == 0:
        gb = B * world
        value = gb * args.steps / (ms * 1e-3)
        line = {"metric": conf["metric"], "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
                "config": {"workload": conf["workload"], "global_batch": gb, "per_gpu_batch": B, "seq_len": L,
                           "parallelism": f"dp{world}", "cuda_graph": graph is not None,
                           "scaling_note": "weak scaling at a fixed per-GPU batch; BASELINE configs[2] (global batch 512) is this line "
                                           "with --batch 512/N and coincides with the default at N = 8",
                           "l2": "no explicit flush: one step touches ~1 GB of activations and weights, 8x the 126 MB L2",
                           "timed_region": "forward + loss + backward" + (" + gradient all-reduce" if world > 1 else "")},
                "clocks": clocks,
                "e2e": {"value": gb * args.steps / (ms_e2e * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": h2d_bytes,
                        "d2h_bytes_per_step": 8, "ms_per_step": ms_e2e / args.steps,
                        "loss_read": "every step's loss is copied to pinned host memory and read there one step late "
                                     "(GraphedPretrainStep.read_loss_async); the last one before the region ends"},
                "gpu_launches": int(launches), "roofline": roofline,
                "model_tflops": flops_per_sample(CFG) * value / 1e12,
                "model_tflops_note": f"FLOPs executed (closed-form attention, {flops_per_sample(CFG) / 1e9:.3f} GF/sample); the reference "
                                     f"executes {flops_per_sample(CFG, True) / 1e9:.3f} GF/sample for the same result",
                "frac_of_sustained_bf16_peak": flops_per_sample(CFG) * value / world / 1e12 / peaks["bf16_tflops_sustained"]}
        if ddp_check is not None:
            line["ddp_check"] = ddp_check
        if comm is not None:
            line.update(comm)
        if adam is not None:
            line["adam_step"] = adam
        if table is not None:
            line["roofline_table"] = table
        if world == 1 and not args.no_gpu_baseline:
            try:
                del graph
                torch.cuda.empty_cache()
                line["gpu_eager_baseline"] = gpu_eager_baseline(CFG, B, dev)
            except Exception as e:
                line["gpu_eager_baseline"] = {"error": f"{type(e).__name__}: {e}"[:300]}
        if world == 1 and not args.no_cpu_baseline:
            base = cpu_reference_baseline(CFG)
            line["cpu_baseline"] = {k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "cpu")}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
