"""Per-operator device time of one eager pretraining step (config 2, bf16): every C-ABI call is bracketed by CUDA
events (ops.PROFILE).  Eager launches are CPU-bound, so each op starts on an idle GPU with the caches in the state the
previous ops left them - unlike an ncu launch list, which flushes L2 before every kernel.

    python scripts/op_profile.py [--steps 5] [--batch 64] > gpurun_out/op_profile.md
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


class AllOps(dict):
    def __contains__(self, k):
        return True

    def __missing__(self, k):
        self[k] = []
        return self[k]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--batch", type=int, default=64)
    args = ap.parse_args()
    import torch
    import proteinbert_b200 as pb
    from proteinbert_b200 import ops
    from bench import CFG
    
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = pb.ProteinBERT(device=dev, precision="bf16", **CFG)
    b = [t.to(dev) for t in pb.data_processing.synthetic_batch(args.batch, CFG["sequences_length"], CFG["num_annotations"], seed=100)]
    X, Y, W = {"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]}
    for _ in range(3):
        pb.pretraining_step(model, X, Y, W, None)
    torch.cuda.synchronize()
    prof = AllOps()
    ops.PROFILE = prof
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        pb.pretraining_step(model, X, Y, W, None)
    e1.record()
    torch.cuda.synchronize()
    ops.PROFILE = None
    rows = []
    for name, evs in prof.items():
        us = [a.elapsed_time(b_) * 1e3 for a, b_ in evs]
        rows.append((sum(us) / args.steps, len(us) / args.steps, sum(us) / len(us), min(us), name))
    rows.sort(reverse=True)
    tot = sum(r[0] for r in rows)
    print(f"# eager step, batch {args.batch}: {tot:.0f} us inside C-ABI calls per step "
          f"(wall {e0.elapsed_time(e1) / args.steps * 1e3:.0f} us per eager step incl. launch gaps)\n")
    print("| C-ABI call | calls/step | us/call (mean) | us/call (min) | us/step | share |")
    print("|---|---:|---:|---:|---:|---:|")
    for t, n, mean, mn, name in rows:
        print(f"| `{name}` | {n:.0f} | {mean:.1f} | {mn:.1f} | {t:.0f} | {100 * t / tot:.1f}% |")
    print("\nper call site (mean us over the steps, in call order within a step):\n")
    for t, n, mean, mn, name in rows:
        n = int(round(n))
        if n < 2:
            continue
        us = [a.elapsed_time(b_) * 1e3 for a, b_ in prof[name]]
        sites = [sum(us[i::n]) / args.steps for i in range(n)]
        print(f"- `{name}`: " + " ".join(f"{v:.0f}" for v in sites))


if __name__ == "__main__":
    main()
