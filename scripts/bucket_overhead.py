"""1 GPU: graph replay time of the step without the gradient bucket, with it (views + one clearing launch), and with the two-graph split."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import proteinbert_b200 as pb
from bench import BASE_CFG as CFG
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = pb.ProteinBERT(device=dev, precision="bf16", **CFG)
b = [t.to(dev) for t in pb.data_processing.synthetic_batch(64, 512, 8943, seed=100)]
X, Y, W = {"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]}
for _ in range(3):
    pb.pretraining_step(model, X, Y, W)
for name, kw in (("plain", {}), ("bucket", {"force_bucket": True}), ("split", {"force_split": True}), ("plain again", {})):
    step = pb.GraphedPretrainStep(model, X, Y, W, warmup=1, **kw)
    for _ in range(5):
        step.replay_local()
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(30):
        step.replay_local()
    e.record(); torch.cuda.synchronize()
    print(f"{name:12s} {a.elapsed_time(e) / 30 * 1e3:8.1f} us per step")
    del step
