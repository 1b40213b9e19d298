"""Target for ncu captures: eager bf16 steps of BASELINE config 2 (B=64, L=512) and nothing else.
    python scripts/ncu_target.py [steps]        # 2 warm-up steps, then `steps` (default 1)"""
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
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
for _ in range(2 + n):
    loss = pb.pretraining_step(model, X, Y, W)
torch.cuda.synchronize()
print("loss", float(loss))
