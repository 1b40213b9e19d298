"""Two (or more) ranks, one process per GPU over NCCL: the gradients GraphedPretrainStep leaves in .grad after its bucket
all-reduce must equal (a) the sum over ranks of the rank-local gradients (ddp_check) and (b) the mean of single-GPU EAGER
gradients of every rank's shard, computed locally on each rank without any exchange -- SURVEY.md 8e's oracle, "the reference
run per shard, gradients averaged".

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 scripts/ddp_check.py --out r.json
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import proteinbert_b200 as pb

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="")
ap.add_argument("--precision", default="fp32")
args = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
cfg = dict(sequences_length=256, num_annotations=517, local_dim=128, global_dim=512, key_dim=64, num_heads=4, num_blocks=4)
torch.manual_seed(1234 + rank)                       # deliberately different seeds: the broadcast must make the ranks agree
model = pb.ProteinBERT(device=dev, precision=args.precision, **cfg)
red = pb.GradAllReducer(model.parameters())
red.broadcast_parameters(0, extra_tensors=model.head_tensors().values())
red.remove()
B = 8
shards = [[t.to(dev) for t in pb.data_processing.synthetic_batch(B, cfg["sequences_length"], cfg["num_annotations"], seed=50 + r)]
          for r in range(world)]
d = lambda b: ({"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]})

# (b) every shard eagerly on this rank, no exchange; mean over shards
want = None
for r in range(world):
    pb.pretraining_step(model, *d(shards[r]))
    g = {n: p.grad.detach().double().clone() for n, p in model.named_parameters()}
    want = g if want is None else {n: want[n] + g[n] for n in g}
want = {n: v / world for n, v in want.items()}

step = pb.GraphedPretrainStep(model, *d(shards[rank]), warmup=1)
check = step.ddp_check(*d(shards[rank]))
step(*d(shards[rank]))
torch.cuda.synchronize()
worst, worst_name, worst_wp = 0.0, "", 0.0
gmax = max(float(v.abs().max()) for v in want.values())
for n, p in model.named_parameters():
    ref = want[n]
    err = float((p.grad.double() - ref).norm())
    den = float(ref.norm()) + 1e-6 * gmax * ref.numel() ** 0.5        # absolute floor for mathematically-zero gradients
    if n == "pretraining_local_output.0.bias":        # mathematically zero (batch-axis softmax): only round-off on either side
        continue
    if n.endswith("W_parameter"):       # a cancelling sum ~1e-4 of the largest gradient: fp32 atomics reorder it from run to run
        worst_wp = max(worst_wp, err / den)
    elif err / den > worst:
        worst, worst_name = err / den, n
t = torch.tensor([check, worst, worst_wp], device=dev, dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
res = {"world": world, "max_rel_err": float(t[0]), "vs_eager_per_shard_mean": float(t[1]), "vs_eager_W_parameter": float(t[2]), "worst_tensor": worst_name,
       "loose_tensors": len(step.bucket.loose), "split": bool(step.split), "precision": args.precision}
if rank == 0:
    print(json.dumps(res))
    if args.out:
        json.dump(res, open(args.out, "w"))
del step
dist.destroy_process_group()
