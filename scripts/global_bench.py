"""Back-to-back timing of the four fused global-track kernels through the C ABI (CUDA events), B=64, G=512, C=128.
    python scripts/global_bench.py [B G C]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import proteinbert_b200 as pb
from proteinbert_b200 import ops

B, G, C = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (64, 512, 128)
dev = torch.device("cuda", 0)
torch.manual_seed(0)
f = lambda *s: torch.randn(*s, device=dev)
g, W1, b1, W3, b3 = f(B, G), f(C, G) * 0.04, f(C), f(G, G) * 0.04, f(G)
zg, g2l, z3, gh = f(B, C), f(B, C), f(B, G), f(B, G)
pool, Wp, ga1, be1, W4, b4, ga2, be2 = f(B, G), f(64), f(G), f(G), f(G, G) * 0.04, f(G), f(G), f(G)
buf = torch.empty(5, B, G, device=dev)
rstd = torch.empty(2, B, device=dev)
st = torch.cuda.current_stream().cuda_stream
p = lambda t: None if t is None else t.data_ptr()


def pre_fwd():
    ops._call("pb_global_pre_fwd", p(g), p(W1), p(b1), p(W3), p(b3), p(zg), p(g2l), p(z3), p(gh), B, G, C, st)


def post_fwd():
    ops._call("pb_global_post_fwd", p(g), p(gh), p(pool), p(Wp), Wp.numel(), p(ga1), p(be1), p(W4), p(b4), p(ga2), p(be2),
              p(buf[0]), p(buf[1]), p(rstd[0]), p(buf[2]), p(buf[3]), p(rstd[1]), p(buf[4]), B, G, st)


dg2 = f(B, G)
out = torch.empty(6, B, G, device=dev)
dWp, vec, dW4 = torch.empty(64, device=dev), torch.empty(5, G, device=dev), torch.empty(G, G, device=dev)


def post_bwd():
    ops._call("pb_global_post_bwd", p(dg2), p(ga2), p(buf[3]), p(rstd[1]), p(buf[2]), p(buf[0]), p(W4), p(ga1), p(buf[1]), p(rstd[0]),
              p(pool), p(Wp), Wp.numel(), p(out[0]), p(out[3]), p(out[4]), p(dWp), p(vec[0]), p(vec[1]), None, p(vec[2]), p(vec[3]),
              p(vec[4]), p(z3), p(out[5]), B, G, st)


dg2l, dg, db1, db3, dzg = f(B, C), torch.empty(B, G, device=dev), torch.empty(C, device=dev), torch.empty(G, device=dev), torch.empty(B, C, device=dev)


def pre_bwd():
    ops._call("pb_global_pre_bwd", p(dg2l), p(out[5]), p(zg), p(z3), p(g), p(W1), p(W3), p(dg), None, p(db1), None, p(db3), B, G, C, 1,
              p(dzg), st)


def matmul_tn():
    ops._call("pb_matmul_tn", p(out[5]), p(g), p(dW4), B, G, G, st)


for name, fn in (("pre_fwd", pre_fwd), ("post_fwd", post_fwd), ("post_bwd", post_bwd), ("pre_bwd", pre_bwd), ("matmul_tn", matmul_tn)):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 100
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    print(f"{name:10s} {a.elapsed_time(b) / n * 1e3:7.2f} us per launch (back to back)")

# ---- cluster units (csrc/gtrack.cu) -----------------------------------------------------------------------------------
if ops.gtrack_ok(B, G, C):
    zg2, z32 = torch.empty(2, B, C, device=dev), torch.empty(2, B, G, device=dev)

    def unit_fwd():
        ops._call("pb_gtrack_fwd", p(g), p(gh), p(pool), p(Wp), Wp.numel(), p(ga1), p(be1), p(W4), p(b4), p(ga2), p(be2), p(buf[0]),
                  p(buf[1]), p(rstd[0]), p(buf[2]), p(buf[3]), p(rstd[1]), p(buf[4]), None, p(W1), p(b1), p(W3), p(b3), p(zg2[0]),
                  p(zg2[1]), p(z32[0]), p(z32[1]), B, G, C, st)

    o2 = torch.empty(8, B, G, device=dev)
    db1_, db3_, dzg_ = torch.empty(C, device=dev), torch.empty(G, device=dev), torch.empty(B, C, device=dev)

    def unit_bwd():
        ops._call("pb_gtrack_bwd", p(dg2l), p(zg2[0]), p(out[5]), None, None, p(dg2), p(W1), p(W3), None, p(dzg_), p(db1_), p(db3_), p(o2[6]),
                  None, p(ga2), p(buf[3]), p(rstd[1]), p(buf[2]), p(W4), p(ga1), p(buf[1]), p(rstd[0]), p(pool), p(Wp), Wp.numel(), p(z3),
                  p(o2[0]), p(o2[1]), p(o2[2]), p(o2[3]), p(o2[5]), p(o2[4]), p(dWp), p(vec[0]), p(vec[1]), p(vec[2]), p(vec[3]), p(vec[4]),
                  B, G, C, st)

    for name, fn in (("unit_fwd", unit_fwd), ("unit_bwd", unit_bwd)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 100
        a.record()
        for _ in range(n):
            fn()
        b.record()
        torch.cuda.synchronize()
        print(f"{name:10s} {a.elapsed_time(b) / n * 1e3:7.2f} us per launch (back to back)")
else:
    print("gtrack: no 16-CTA cluster available / shape not covered")

import ctypes
if ops.gtrack_ok(B, G, C):
    dbg = torch.zeros(64, dtype=torch.int64, device=dev)
    lib = pb._lib.lib()
    lib.pb_debug_buffer.argtypes = [ctypes.c_void_p]
    lib.pb_debug_buffer(dbg.data_ptr())
    unit_fwd()
    torch.cuda.synchronize()
    lib.pb_debug_buffer(None)
    t = dbg.cpu().tolist()
    if t[13]:
        names = ["start", "init+W issue+cluster sync+pdl", "P0 rows LN1", "cluster sync", "W slice arrived", "A tile arrived", "mma", "reduce+epilogue",
                 "cluster sync", "P2 rows LN2", "cluster sync", "W+A arrived (P3)", "mma (P3)", "reduce+epilogue (P3)"]
        for i in range(1, 14):
            print(f"  fwd {names[i]:34s} +{t[i] - t[i - 1]:6d} ns  (t = {t[i] - t[0]})")
