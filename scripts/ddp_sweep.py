"""N ranks: time the data-parallel step under several exchange shapes in ONE process group (a fresh torchrun per shape costs
a minute of start-up each).  Each shape = env overrides read by FlatGradBucket / GraphedPretrainStep at construction.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29653 scripts/ddp_sweep.py [out.json]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import proteinbert_b200 as pb
from bench import BASE_CFG as CFG

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
torch.manual_seed(0)
model = pb.ProteinBERT(device=dev, precision="bf16", **CFG)
b = [t.to(dev) for t in pb.data_processing.synthetic_batch(64, 512, 8943, seed=100 + rank)]
X, Y, W = {"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]}
for _ in range(3):
    pb.pretraining_step(model, X, Y, W)

KEYS = ("PB_B200_EXCHANGE", "PB_B200_DDP_MODE", "PB_B200_DDP_CUT", "PB_B200_DDP_JOIN_AT_CUT", "PB_B200_NVLS_CTAS", "PB_B200_NVLS_THREADS",
        "PB_B200_NVLS_EARLY_CTAS", "PB_B200_NVLS_EARLY_THREADS", "PB_B200_NVLS_EARLY_INFLIGHT")
ex = "nvls" if world >= 4 else os.environ.get("SWEEP_EXCHANGE", "nvls")
SHAPES = [{'name': 'single (no overlap)', 'PB_B200_DDP_MODE': 'single'},
    {'name': 'event cut3 16x512x8', 'PB_B200_DDP_CUT': '3'},
    {'name': 'event cut3 early 8x256x4', 'PB_B200_DDP_CUT': '3', 'PB_B200_NVLS_EARLY_CTAS': '8', 'PB_B200_NVLS_EARLY_THREADS': '256', 'PB_B200_NVLS_EARLY_INFLIGHT': '4'},
    {'name': 'event cut3 early 32x128x2', 'PB_B200_DDP_CUT': '3', 'PB_B200_NVLS_EARLY_CTAS': '32', 'PB_B200_NVLS_EARLY_THREADS': '128', 'PB_B200_NVLS_EARLY_INFLIGHT': '2'},
    {'name': 'event cut3 early 148x32x2', 'PB_B200_DDP_CUT': '3', 'PB_B200_NVLS_EARLY_CTAS': '148', 'PB_B200_NVLS_EARLY_THREADS': '32', 'PB_B200_NVLS_EARLY_INFLIGHT': '2'},
    {'name': 'event cut3 early 16x128x4', 'PB_B200_DDP_CUT': '3', 'PB_B200_NVLS_EARLY_CTAS': '16', 'PB_B200_NVLS_EARLY_THREADS': '128', 'PB_B200_NVLS_EARLY_INFLIGHT': '4'},
    {'name': 'event cut3,1 early 8x256x4', 'PB_B200_DDP_CUT': '3,1', 'PB_B200_NVLS_EARLY_CTAS': '8', 'PB_B200_NVLS_EARLY_THREADS': '256', 'PB_B200_NVLS_EARLY_INFLIGHT': '4'},
    {'name': 'event cut3,1 early 32x128x2', 'PB_B200_DDP_CUT': '3,1', 'PB_B200_NVLS_EARLY_CTAS': '32', 'PB_B200_NVLS_EARLY_THREADS': '128', 'PB_B200_NVLS_EARLY_INFLIGHT': '2'},
    {'name': 'event cut3,1 early 148x32x2', 'PB_B200_DDP_CUT': '3,1', 'PB_B200_NVLS_EARLY_CTAS': '148', 'PB_B200_NVLS_EARLY_THREADS': '32', 'PB_B200_NVLS_EARLY_INFLIGHT': '2'},
    {'name': 'event cut3,1 early 16x128x4', 'PB_B200_DDP_CUT': '3,1', 'PB_B200_NVLS_EARLY_CTAS': '16', 'PB_B200_NVLS_EARLY_THREADS': '128', 'PB_B200_NVLS_EARLY_INFLIGHT': '4'},
    {'name': 'event cut4,2,1 early 8x256x4', 'PB_B200_DDP_CUT': '4,2,1', 'PB_B200_NVLS_EARLY_CTAS': '8', 'PB_B200_NVLS_EARLY_THREADS': '256', 'PB_B200_NVLS_EARLY_INFLIGHT': '4'},
    {'name': 'event cut4,2,1 early 32x128x2', 'PB_B200_DDP_CUT': '4,2,1', 'PB_B200_NVLS_EARLY_CTAS': '32', 'PB_B200_NVLS_EARLY_THREADS': '128', 'PB_B200_NVLS_EARLY_INFLIGHT': '2'},
    {'name': 'event cut4,2,1 early 148x32x2', 'PB_B200_DDP_CUT': '4,2,1', 'PB_B200_NVLS_EARLY_CTAS': '148', 'PB_B200_NVLS_EARLY_THREADS': '32', 'PB_B200_NVLS_EARLY_INFLIGHT': '2'},
    {'name': 'event cut4,2,1 early 16x128x4', 'PB_B200_DDP_CUT': '4,2,1', 'PB_B200_NVLS_EARLY_CTAS': '16', 'PB_B200_NVLS_EARLY_THREADS': '128', 'PB_B200_NVLS_EARLY_INFLIGHT': '4'},
    {'name': 'event cut3,1 early 8x256x4 final 32x512x8', 'PB_B200_DDP_CUT': '3,1', 'PB_B200_NVLS_EARLY_CTAS': '8', 'PB_B200_NVLS_EARLY_THREADS': '256', 'PB_B200_NVLS_EARLY_INFLIGHT': '4', 'PB_B200_NVLS_CTAS': '32'},
    {'name': 'event cut3,1 early 8x256x4 final 8x512x8', 'PB_B200_DDP_CUT': '3,1', 'PB_B200_NVLS_EARLY_CTAS': '8', 'PB_B200_NVLS_EARLY_THREADS': '256', 'PB_B200_NVLS_EARLY_INFLIGHT': '4', 'PB_B200_NVLS_CTAS': '8'}]
def _sh(c, t, f, cut="3,1"):
    return dict(name=f"nvls event cut{cut} early {c}x{t}x{f}", PB_B200_EXCHANGE="nvls", PB_B200_DDP_CUT=cut, PB_B200_NVLS_EARLY_CTAS=str(c),
                PB_B200_NVLS_EARLY_THREADS=str(t), PB_B200_NVLS_EARLY_INFLIGHT=str(f))


if os.environ.get("SWEEP_SET") == "small_world":          # 2 GPUs: NCCL against the NVLS kernel at several intensities
    SHAPES = [dict(name="nccl single", PB_B200_EXCHANGE="nccl", PB_B200_DDP_MODE="single"),
              dict(name="nccl event cut3", PB_B200_EXCHANGE="nccl", PB_B200_DDP_CUT="3"),
              dict(name="nccl event cut3,1", PB_B200_EXCHANGE="nccl", PB_B200_DDP_CUT="3,1"),
              dict(name="nccl event cut4,2,1", PB_B200_EXCHANGE="nccl", PB_B200_DDP_CUT="4,2,1"),
              dict(name="nvls single", PB_B200_EXCHANGE="nvls", PB_B200_DDP_MODE="single"),
              _sh(16, 128, 4), _sh(32, 128, 4), _sh(32, 256, 4), _sh(64, 128, 4), _sh(64, 128, 8), _sh(32, 128, 4, "4,2,1"), _sh(64, 128, 4, "4,2,1")]
if os.environ.get("SWEEP_SET") == "mid_world":            # 4 GPUs
    SHAPES = [dict(name="nccl event cut3,1", PB_B200_EXCHANGE="nccl", PB_B200_DDP_CUT="3,1"),
              dict(name="nvls single", PB_B200_EXCHANGE="nvls", PB_B200_DDP_MODE="single"),
              _sh(16, 128, 4), _sh(32, 128, 4), _sh(64, 128, 4), _sh(16, 256, 4)]
only = os.environ.get("SWEEP_ONLY")
if only:
    SHAPES = [s for i, s in enumerate(SHAPES) if str(i) in only.split(",")]


def timed(fn, steps):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize(dev)
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    e.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([a.elapsed_time(e) / steps * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


rows = []
for shape in SHAPES:
    for k in KEYS:
        os.environ.pop(k, None)
    os.environ["PB_B200_EXCHANGE"] = ex
    for k, v in shape.items():
        if k != "name":
            os.environ[k] = v
    step = pb.GraphedPretrainStep(model, X, Y, W, warmup=2)
    trials = sorted(timed(lambda: step(X, Y, W), 20) for _ in range(4))
    step_us, step2_us = trials[0], trials[-1]          # best and worst of four 20-step runs
    local_us = timed(step.replay_local, 20)
    chk = step.ddp_check(X, Y, W)
    row = dict(name=shape["name"], mode=step.mode, exchange=step.bucket.exchange, step_us=round(step_us, 1), step_worst_us=round(step2_us, 1), step_median_us=round((trials[1] + trials[2]) / 2, 1),
               compute_only_us=round(local_us, 1), cut_bytes=[4 * o for o in (step.cut_offsets or [])], ddp_check=chk)
    rows.append(row)
    if rank == 0:
        print(json.dumps(row), flush=True)
    del step
    torch.cuda.synchronize(dev)
if rank == 0 and len(sys.argv) > 1:
    json.dump(rows, open(sys.argv[1], "w"), indent=1)
dist.destroy_process_group()
