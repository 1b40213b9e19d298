"""torchrun probe: cost of the flat-bucket gradient exchange (gather, NCCL all-reduce, scatter) at the model's size."""
import os
import sys
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n = 16175241
for dtype in (torch.float32, torch.bfloat16):
    flat = torch.randn(n, device="cuda").to(dtype)
    for _ in range(5):
        dist.all_reduce(flat)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier(); torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        dist.all_reduce(flat)
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        ms = e0.elapsed_time(e1) / 20
        print(f"all_reduce {dtype} {n} elems, world {world}: {ms:.3f} ms, busbw {2 * (world - 1) / world * flat.numel() * flat.element_size() / ms / 1e6:.1f} GB/s", flush=True)
# chunked: 4 buckets back to back
flat = torch.randn(n, device="cuda")
chunks = list(flat.chunk(4))
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    for c in chunks:
        dist.all_reduce(c)
e1.record(); torch.cuda.synchronize()
if rank == 0:
    print(f"4 buckets: {e0.elapsed_time(e1) / 20:.3f} ms", flush=True)
dist.destroy_process_group()
