"""Phase timestamps (globaltimer, CTA 0) of the fused global-track backward kernels, taken in a real step.
    python scripts/global_phase_probe.py"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import proteinbert_b200 as pb
from bench import BASE_CFG as CFG

dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = pb.ProteinBERT(device=dev, precision="bf16", **CFG)
b = [t.to(dev) for t in pb.data_processing.synthetic_batch(64, CFG["sequences_length"], CFG["num_annotations"], seed=100)]
X, Y, W = {"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]}
for _ in range(3):
    pb.pretraining_step(model, X, Y, W, None)
dbg = torch.zeros(64, dtype=torch.int64, device=dev)
lib = pb._lib.lib()
lib.pb_debug_buffer.argtypes = [ctypes.c_void_p]
lib.pb_debug_buffer(dbg.data_ptr())
pb.pretraining_step(model, X, Y, W, None)
torch.cuda.synchronize()
lib.pb_debug_buffer(None)
t = dbg.cpu().tolist()
names = {0: "start", 1: "LN2 bwd rows done", 2: "after grid.sync", 3: "A tile (dz4) staged", 4: "D slice + column sums (8 threads)",
         5: "dW accumulate", 6: "dz4 W4 product", 7: "stores", 8: "after grid.sync", 9: "LN1 bwd rows", 10: "dgamma1/dbeta1 (8 threads)"}
print("post_bwd (last launch of the step), ns since kernel start of CTA 0:")
for i in range(1, 11):
    print(f"  {names[i]:40s} +{t[32 + i] - t[31 + i]:6d}   (t = {t[32 + i] - t[32]})")
names = {16: "start", 17: "A tile [dz3|dzg] staged (gelu')", 18: "D slice + column sums", 19: "dW accumulate", 20: "dz W product", 21: "stores"}
print(f"pre_bwd: weight slices loaded +{t[22] - t[16]}, dz3 tile staged +{t[23] - t[22]}, dzg tile (gelu') +{t[17] - t[23]}")
for i in range(17, 22):
    print(f"  {names[i]:40s} +{t[i] - t[i - 1]:6d}   (t = {t[i] - t[16]})")

print("post_fwd (last launch of the step):")
for a, b_, what in ((48, 49, "W4 slice loaded"), (49, 50, "LN1 rows"), (50, 51, "grid.sync"), (51, 52, "g1 tile staged (TMA)"), (52, 53, "mma product"),
                    (53, 54, "z4 stores"), (54, 55, "grid.sync"), (55, 56, "LN2 rows")):
    print(f"  {what:40s} +{t[b_] - t[a]:6d}   (t = {t[b_] - t[48]})")
