"""torchrun probe: where the data-parallel step's extra time goes (replay / gather+scatter / all-reduce)."""
import os
import sys
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import proteinbert_b200 as pb
from oracle import model_oracle

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
CFG = dict(sequences_length=512, num_annotations=8943, local_dim=128, global_dim=512, key_dim=64, num_heads=4, num_blocks=6)
torch.manual_seed(0)
model = pb.ProteinBERT(device=dev, precision="bf16", **CFG)
b = [t.to(dev) for t in model_oracle.synthetic_batch(64, 512, 8943, seed=rank)]
X, Y, W = {"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]}
t = torch.ones(1, device=dev); dist.all_reduce(t)
step = pb.GraphedPretrainStep(model, X, Y, W)
bk = step.bucket


def timeit(fn, n=30):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


res = {
    "replay": timeit(lambda: step.graph.replay()),
    "replay+gather+scatter": timeit(lambda: (step.graph.replay(), bk.gather(), bk.scatter())),
    "replay+allreduce": timeit(lambda: (step.graph.replay(), bk.all_reduce())),
    "full": timeit(lambda: step(X, Y, W)),
    "allreduce only": timeit(lambda: bk.all_reduce()),
    "gather+scatter only": timeit(lambda: (bk.gather(), bk.scatter())),
}
if rank == 0:
    print({k: round(v, 3) for k, v in res.items()}, flush=True)
torch.cuda.synchronize()
os._exit(0)
