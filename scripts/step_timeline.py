"""Timeline of ONE graphed step (CUPTI activity records of a graph replay): per kernel start offset, duration, stream.
    python scripts/step_timeline.py [out.csv]"""
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
import proteinbert_b200 as pb
from bench import BASE_CFG as CFG

dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = pb.ProteinBERT(device=dev, precision="bf16", **CFG)
b = [t.to(dev) for t in pb.data_processing.synthetic_batch(64, CFG["sequences_length"], CFG["num_annotations"], seed=100)]
X, Y, W = {"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]}
for _ in range(3):
    pb.pretraining_step(model, X, Y, W, None)
step = pb.GraphedPretrainStep(model, X, Y, W, warmup=1)
for _ in range(3):
    step.graph.replay()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step.graph.replay()
    torch.cuda.synchronize()
with tempfile.TemporaryDirectory() as td:
    path = os.path.join(td, "t.json")
    prof.export_chrome_trace(path)
    tr = json.load(open(path))
ev = sorted([e for e in tr["traceEvents"] if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy") and "dur" in e], key=lambda e: e["ts"])
t0 = ev[0]["ts"]
streams = {}
rows = []
for e in ev:
    sid = e.get("args", {}).get("stream", e.get("tid"))
    streams.setdefault(sid, len(streams))
    name = e["name"]
    for pre in ("void ", "(anonymous namespace)::", "<unnamed>::"):
        name = name.replace(pre, "")
    rows.append((e["ts"] - t0, e["dur"], streams[sid], name[:70]))
out = sys.argv[1] if len(sys.argv) > 1 else None
lines = ["start_us,dur_us,stream,kernel"] + [f"{a:.1f},{d:.1f},{s},{n}" for a, d, s, n in rows]
if out:
    open(out, "w").write("\n".join(lines) + "\n")
end = max(a + d for a, d, _, _ in rows)
print(f"step span {end:.1f} us, {len(rows)} kernels, streams {len(streams)}")
# main stream = stream of the first kernel; gaps on it
main = rows[0][2]
busy = sum(d for a, d, s, n in rows if s == main)
print(f"main stream busy {busy:.1f} us, idle {end - busy:.1f} us")
prev_end = 0.0
for a, d, s, n in rows:
    if s != main:
        continue
    gap = a - prev_end
    print(f"{a:8.1f} {d:6.1f} gap {gap:5.1f}  {n}")
    prev_end = a + d
