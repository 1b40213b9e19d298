"""torchrun probe: where the data-parallel step's extra time goes (replay / gather+scatter / all-reduce)."""
import os
import sys
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import proteinbert_b200 as pb

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
_opts = dist.ProcessGroupNCCL.Options()
_opts.is_high_priority_stream = True
dist.init_process_group("nccl", device_id=dev, pg_options=_opts)
CFG = dict(sequences_length=512, num_annotations=8943, local_dim=128, global_dim=512, key_dim=64, num_heads=4, num_blocks=6)
torch.manual_seed(0)
model = pb.ProteinBERT(device=dev, precision="bf16", **CFG)
b = [t.to(dev) for t in pb.data_processing.synthetic_batch(64, 512, 8943, seed=rank)]
X, Y, W = {"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]}
t = torch.ones(1, device=dev); dist.all_reduce(t)


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


res = {}
plain = pb.GraphedPretrainStep(model, X, Y, W, overlap=False)
bk = plain.bucket
res["replay only"] = timeit(lambda: plain.graph.replay())
res["one graph + bucket after"] = timeit(lambda: plain(X, Y, W))
res["allreduce only"] = timeit(lambda: bk.all_reduce())
res["gather+scatter only"] = timeit(lambda: (bk.gather(), bk.scatter()))
plain(X, Y, W)            # a clean step: the timing loops above rescaled the gradients
torch.cuda.synchronize()
g_plain = {n: p.grad.clone() for n, p in model.named_parameters()}
del plain
split = pb.GraphedPretrainStep(model, X, Y, W, overlap=True)
res["two graphs, overlapped"] = timeit(lambda: split(X, Y, W))
worst = max(float((p.grad - g_plain[n]).norm() / g_plain[n].norm()) for n, p in model.named_parameters() if float(g_plain[n].abs().max()) > 1e-6)
if rank == 0:
    print({k: round(v, 3) for k, v in res.items()}, "max rel grad diff split vs plain: %.2e" % worst, flush=True)
torch.cuda.synchronize()
os._exit(0)
