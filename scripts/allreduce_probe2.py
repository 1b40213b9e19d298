"""N ranks: NCCL all-reduce vs the NVLS (multimem) exchange on a symmetric flat buffer of the gradient bucket's size.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 scripts/allreduce_probe2.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm
import proteinbert_b200 as pb
from proteinbert_b200 import ops

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = 16175360
plain = torch.randn(n, device=dev)
buf = symm.empty(n, dtype=torch.float32, device=dev)
hdl = symm.rendezvous(buf, dist.group.WORLD.group_name)
print(rank, "multicast_ptr", hex(hdl.multicast_ptr) if hdl.multicast_ptr else None, flush=True)


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record(); torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


st = lambda: torch.cuda.current_stream().cuda_stream
res = {"nccl": timed(lambda: dist.all_reduce(plain))}
gname = dist.group.WORLD.group_name
for name in ("multimem_all_reduce_", "two_shot_all_reduce_"):
    try:
        op = getattr(torch.ops.symm_mem, name)
        res["torch." + name] = timed(lambda: op(buf, "sum", gname))
    except Exception as e:
        res["torch." + name] = f"{type(e).__name__}: {e}"[:120]
if hdl.multicast_ptr:
    for ctas in (8, 16, 32, 64, 128):
        def own():
            hdl.barrier(channel=0)
            ops._call("pb_multimem_allreduce", hdl.multicast_ptr, 0, n, rank, world, ctas, st())
            hdl.barrier(channel=1)
        res[f"own multimem kernel, {ctas} CTAs (+2 barriers)"] = timed(own)
    # correctness
    buf.copy_(torch.full((n,), float(rank + 1), device=dev) + torch.arange(n, device=dev) % 7)
    torch.cuda.synchronize(); dist.barrier()
    hdl.barrier(channel=0)
    ops._call("pb_multimem_allreduce", hdl.multicast_ptr, 0, n, rank, world, 32, st())
    hdl.barrier(channel=1)
    torch.cuda.synchronize()
    want = world * (world + 1) / 2 + world * (torch.arange(n, device=dev) % 7)
    res["own kernel max abs err"] = float((buf - want).abs().max())
if rank == 0:
    for k, v in res.items():
        print(f"{k:50s} {v}")
dist.destroy_process_group()
