"""Cycle breakdown of the halo conv kernel (CTA 0): MMA thread waits, epilogue time."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import proteinbert_b200 as pb
from proteinbert_b200 import ops
lib = pb._lib.lib()
lib.pb_debug_buffer.argtypes = [ctypes.c_void_p]
B, L, C, KS, DIL = 64, 512, 128, 9, 5
g = torch.Generator().manual_seed(0)
x = torch.randn(B, L, C, generator=g).cuda().bfloat16()
wn, ww = (torch.randn(C, C, KS, generator=g) * 0.05).cuda(), (torch.randn(C, C, KS, generator=g) * 0.05).cuda()
bn, bw, g2l = torch.randn(C).cuda(), torch.randn(C).cuda(), torch.randn(B, C).cuda()
wn_f, wn_b = ops.conv_pack(wn, 1); ww_f, ww_b = ops.conv_pack(ww, 1)
zn, zw, u = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
stats = torch.zeros(B, 2, dtype=torch.float64, device="cuda")
dbg = torch.zeros(8, dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def fwd():
    ops._call("pb_conv_pair_fwd", x.data_ptr(), wn_f.data_ptr(), ww_f.data_ptr(), bn.data_ptr(), bw.data_ptr(), g2l.data_ptr(),
              zn.data_ptr(), zw.data_ptr(), u.data_ptr(), stats.data_ptr(), B, L, C, KS, DIL, 1, st)
for _ in range(3): fwd()
lib.pb_debug_buffer(dbg.data_ptr())
fwd(); torch.cuda.synchronize()
d = dbg.cpu().tolist()
print("fwd: mma issue loop %d cyc (wait weights %d, wait halo %d), all MMAs done at %d; epilogue waited %d, epilogue work %d" % tuple(d[:6]))
lib.pb_debug_buffer(None)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): fwd()
e1.record(); torch.cuda.synchronize()
print("fwd avg us", e0.elapsed_time(e1) / 20 * 1000)

# ---- data-gradient variant ----
du = torch.randn(B, L, C, generator=g).cuda().bfloat16()
dzn, dzw = torch.randn(B, L, C, generator=g).cuda().bfloat16(), torch.randn(B, L, C, generator=g).cuda().bfloat16()
dx = torch.empty_like(x)
def dgrad():
    ops._call("pb_conv_pair_bwd", du.data_ptr(), None, None, x.data_ptr(), wn_b.data_ptr(), ww_b.data_ptr(), dzn.data_ptr(), dzw.data_ptr(),
              dx.data_ptr(), None, None, None, None, None, B, L, C, KS, DIL, 1, None, 0, st)
for _ in range(3): dgrad()
dbg.zero_()
lib.pb_debug_buffer(dbg.data_ptr())
dgrad(); torch.cuda.synchronize()
d = dbg.cpu().tolist()
print("dgrad: mma issue loop %d cyc (wait weights %d, wait halo %d), all MMAs done at %d; epilogue waited %d, epilogue work %d" % tuple(d[:6]))
lib.pb_debug_buffer(None)
e0.record()
for _ in range(20): dgrad()
e1.record(); torch.cuda.synchronize()
print("dgrad avg us", e0.elapsed_time(e1) / 20 * 1000)
