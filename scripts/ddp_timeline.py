"""N ranks: CUPTI timeline of ONE data-parallel step on rank 0 (graph kernels, exchange kernels, barriers) to see where the
exchange lands relative to backward.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29651 scripts/ddp_timeline.py out.csv"""
import json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from torch.profiler import ProfilerActivity, profile
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
step = pb.GraphedPretrainStep(model, X, Y, W, warmup=1)
for _ in range(5):
    step(X, Y, W)
dist.barrier(); torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        step(X, Y, W)
    torch.cuda.synchronize()
if rank == 0:
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "t.json")
        prof.export_chrome_trace(path)
        tr = json.load(open(path))
    ev = sorted([e for e in tr["traceEvents"] if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy") and "dur" in e], key=lambda e: e["ts"])
    t0 = ev[0]["ts"]
    streams = {}
    lines = ["start_us,dur_us,stream,kernel"]
    for e in ev:
        sid = e.get("args", {}).get("stream", e.get("tid"))
        streams.setdefault(sid, len(streams))
        name = e["name"].replace("void ", "").replace("(anonymous namespace)::", "")[:60]
        lines.append(f"{e['ts'] - t0:.1f},{e['dur']:.1f},{streams[sid]},{name}")
    open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/ddp_timeline.csv", "w").write("\n".join(lines) + "\n")
    print("mode", step.mode, "exchange", step.bucket.exchange, "kernels", len(ev), "streams", len(streams))
dist.destroy_process_group()
