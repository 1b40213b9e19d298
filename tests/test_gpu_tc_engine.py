"""GPU: the tcgen05 tensor-core engine against the fp32-accumulating SIMT engine on identical bf16 operands,
op by op through the C ABI (pb_set_tensor_cores toggles the engine).  Both engines round the same inputs to
bf16 and accumulate in fp32, so they may differ only by summation order and by one-ulp flips of bf16 outputs:
rel-Frobenius <= 3e-3 for bf16 outputs, <= 1e-3 for fp32 outputs."""
import pytest
import torch

from helpers import rel_fro

pytestmark = pytest.mark.gpu

B, L, C, G, KS, DIL = 3, 256, 128, 512, 9, 5


@pytest.fixture(scope="module")
def pb():
    import proteinbert_b200
    return proteinbert_b200


def both_engines(pb, fn, modes=(0, 1), op=None):
    """Run fn on the SIMT engine and on the tensor-core engine.  `op` names the C-ABI op family whose engine choice is
    asserted first (pb_tc_path), so that a shape which silently fell back to SIMT cannot compare SIMT with SIMT."""
    lib = pb._lib.lib()
    if op is not None:
        for mode in modes:
            lib.pb_set_tensor_cores(mode)
            try:
                path = lib.pb_tc_path(op.encode(), B, L, C, G, KS, DIL)
            finally:
                lib.pb_set_tensor_cores(1)
            want = {0: 0, 1: 2 if op != "local_linear" else 1, 4: 1}[mode]
            assert path == want, f"{op}: engine switch {mode} takes path {path}, expected {want}"
    outs = []
    for enabled in modes:
        lib.pb_set_tensor_cores(enabled)
        try:
            outs.append([o.float().clone() for o in fn()])
            torch.cuda.synchronize()
        finally:
            lib.pb_set_tensor_cores(1)
    return outs


def compare(names, simt, tc, tol_bf16=3e-3):
    bad = []
    for n, a, b in zip(names, simt, tc):
        e = rel_fro(b, a)
        print(f"  {n}: rel_fro {e:.3e} |ref| {float(a.norm()):.3e}")
        if not e <= tol_bf16:
            bad.append((n, e))
    assert not bad, bad


def wsargs(pb):
    from proteinbert_b200 import ops
    return ops._wgrad_ws(torch.device("cuda", torch.cuda.current_device()))


def rnd(*shape, scale=1.0, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale)


def test_conv_pair_forward(pb):
    from proteinbert_b200 import ops
    x = rnd(B, L, C, seed=1).cuda().bfloat16()
    wn, ww = rnd(C, C, KS, scale=0.05, seed=2).cuda(), rnd(C, C, KS, scale=0.05, seed=3).cuda()
    bn, bw, g2l = rnd(C, seed=4).cuda(), rnd(C, seed=5).cuda(), rnd(B, C, seed=6).cuda()
    wn_f, _ = ops.conv_pack(wn, 1)
    ww_f, _ = ops.conv_pack(ww, 1)

    def run():
        zn, zw, u = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
        stats = torch.zeros(B, 2, dtype=torch.float64, device="cuda")
        ops._call("pb_conv_pair_fwd", x.data_ptr(), wn_f.data_ptr(), ww_f.data_ptr(), bn.data_ptr(), bw.data_ptr(), g2l.data_ptr(),
                  zn.data_ptr(), zw.data_ptr(), u.data_ptr(), stats.data_ptr(), B, L, C, KS, DIL, 1, torch.cuda.current_stream().cuda_stream)
        return zn, zw, u, stats

    simt, tc = both_engines(pb, run, op="conv")
    compare(["zn", "zw", "u", "stats"], simt, tc)
    # and against torch conv1d in fp64 on the bf16-rounded operands (edge rows exercise the zero padding)
    xr = x.double().cpu().permute(0, 2, 1)
    ref = torch.nn.functional.conv1d(xr, ww.bfloat16().double().cpu(), bw.double().cpu(), padding="same", dilation=DIL).permute(0, 2, 1)
    assert rel_fro(tc[1], ref) < 4e-3
    assert rel_fro(tc[1][:, :24], ref[:, :24]) < 4e-3 and rel_fro(tc[1][:, -24:], ref[:, -24:]) < 4e-3


def test_conv_pair_backward(pb):
    from proteinbert_b200 import ops
    x = rnd(B, L, C, seed=1).cuda().bfloat16()
    zn, zw = rnd(B, L, C, seed=7).cuda().bfloat16(), rnd(B, L, C, seed=8).cuda().bfloat16()
    du = rnd(B, L, C, scale=0.1, seed=9).cuda().bfloat16()
    wn, ww = rnd(C, C, KS, scale=0.05, seed=2).cuda(), rnd(C, C, KS, scale=0.05, seed=3).cuda()
    _, wn_b = ops.conv_pack(wn, 1)
    _, ww_b = ops.conv_pack(ww, 1)

    def run():
        dzn, dzw, dx = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
        dWn, dWw = torch.empty(C, C, KS, device="cuda"), torch.empty(C, C, KS, device="cuda")
        dbn, dbw, dg2l = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda"), torch.zeros(B, C, device="cuda")   # accumulated outputs
        ops._call("pb_conv_pair_bwd", du.data_ptr(), zn.data_ptr(), zw.data_ptr(), x.data_ptr(), wn_b.data_ptr(), ww_b.data_ptr(),
                  dzn.data_ptr(), dzw.data_ptr(), dx.data_ptr(), dWn.data_ptr(), dWw.data_ptr(), dbn.data_ptr(), dbw.data_ptr(),
                  dg2l.data_ptr(), B, L, C, KS, DIL, 1, *wsargs(pb), torch.cuda.current_stream().cuda_stream)
        return dx, dWn, dWw, dbn, dbw, dg2l

    simt, tc = both_engines(pb, run, op="conv")
    compare(["dx", "dWn", "dWw", "dbn", "dbw", "dg2l"], simt, tc)


def test_local_linear(pb):
    from proteinbert_b200 import ops
    h = rnd(B, L, C, seed=1).cuda().bfloat16()
    W = rnd(C, C, scale=0.1, seed=2).cuda()
    bias = rnd(C, seed=3).cuda()
    Ww, WTw = ops.to_wt(W, 1), ops.to_wt(W.t(), 1)
    du2 = rnd(B, L, C, scale=0.1, seed=4).cuda().bfloat16()

    def run():
        z, u2 = torch.empty_like(h), torch.empty_like(h)
        stats = torch.zeros(B, 2, dtype=torch.float64, device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        ops._call("pb_local_linear_fwd", h.data_ptr(), Ww.data_ptr(), bias.data_ptr(), z.data_ptr(), u2.data_ptr(), stats.data_ptr(), B, L, C, 1, st)
        dz, dh = torch.empty_like(h), torch.empty_like(h)
        dW, db = torch.empty(C, C, device="cuda"), torch.zeros(C, device="cuda")
        ops._call("pb_local_linear_bwd", du2.data_ptr(), z.data_ptr(), h.data_ptr(), Ww.data_ptr(), WTw.data_ptr(), dz.data_ptr(),
                  dh.data_ptr(), dW.data_ptr(), db.data_ptr(), B, L, C, 1, *wsargs(pb), st)
        return z, u2, stats, dh, dW, db

    simt, tc = both_engines(pb, run, op="local_linear")
    compare(["z", "u2", "stats", "dh", "dW", "db"], simt, tc)


def test_attention_pool(pb):
    from proteinbert_b200 import ops
    h = rnd(B, L, C, seed=1).cuda().bfloat16()
    wv = rnd(C, G, scale=0.3, seed=2).cuda()
    wvTw, wvw = ops.to_wt(wv.t(), 1), ops.to_wt(wv, 1)
    dpool = rnd(B, G, scale=0.1, seed=3).cuda()
    dh_in = rnd(B, L, C, scale=0.1, seed=4).cuda().bfloat16()

    def run():
        st = torch.cuda.current_stream().cuda_stream
        pool = torch.zeros(B, G, device="cuda")
        ops._call("pb_attn_pool_fwd", h.data_ptr(), wvTw.data_ptr(), pool.data_ptr(), B, L, C, G, 1, st)
        dzv = torch.empty(B * L, G, device="cuda", dtype=torch.bfloat16)
        dh = torch.empty_like(h)
        dWvT = torch.empty(G, C, device="cuda")
        ops._call("pb_attn_pool_bwd", h.data_ptr(), wvTw.data_ptr(), wvw.data_ptr(), dpool.data_ptr(), dh_in.data_ptr(), dzv.data_ptr(),
                  dh.data_ptr(), dWvT.data_ptr(), 0, B, L, C, G, 1, *wsargs(pb), st)
        return pool, dzv, dh, dWvT

    simt, tc = both_engines(pb, run, op="attn_pool")
    compare(["pool", "dzv", "dh", "dWvT"], simt, tc)
    ref = (torch.nn.functional.gelu(h.double().cpu() @ wvw.double().cpu())).sum(1)
    assert rel_fro(tc[0], ref) < 2e-3


def test_ln_apply_fast_path(pb):
    from proteinbert_b200 import ops
    u = rnd(B, L, C, seed=1).cuda().bfloat16()
    gamma, beta = (1 + 0.2 * rnd(L, C, seed=2)).cuda(), (0.1 * rnd(L, C, seed=3)).cuda()
    stats = torch.stack([u.double().sum((1, 2)), (u.double() ** 2).sum((1, 2))], 1).contiguous()
    h = torch.empty_like(u)
    ops._call("pb_ln_joint_apply", u.data_ptr(), stats.data_ptr(), gamma.data_ptr(), beta.data_ptr(), h.data_ptr(), B, L, C, 1e-5, 1,
              torch.cuda.current_stream().cuda_stream)
    ref = torch.nn.functional.layer_norm(u.double(), (L, C), gamma.double(), beta.double(), 1e-5)
    assert rel_fro(h, ref) < 3e-3


@pytest.mark.parametrize("mode", [1, 4])
def test_conv_pair_both_kernel_generations(pb, mode):
    """The 256-position halo kernel (tc_conv2.cu, mode 1 = default) and the first-generation kernel (tc_conv.cu,
    mode 4) against the SIMT engine."""
    from proteinbert_b200 import ops
    x = rnd(B, L, C, seed=1).cuda().bfloat16()
    wn, ww = rnd(C, C, KS, scale=0.05, seed=2).cuda(), rnd(C, C, KS, scale=0.05, seed=3).cuda()
    bn, bw, g2l = rnd(C, seed=4).cuda(), rnd(C, seed=5).cuda(), rnd(B, C, seed=6).cuda()
    wn_f, wn_b = ops.conv_pack(wn, 1)
    ww_f, ww_b = ops.conv_pack(ww, 1)
    du = rnd(B, L, C, scale=0.1, seed=9).cuda().bfloat16()

    def run():
        st = torch.cuda.current_stream().cuda_stream
        zn, zw, u = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
        stats = torch.zeros(B, 2, dtype=torch.float64, device="cuda")
        ops._call("pb_conv_pair_fwd", x.data_ptr(), wn_f.data_ptr(), ww_f.data_ptr(), bn.data_ptr(), bw.data_ptr(), g2l.data_ptr(),
                  zn.data_ptr(), zw.data_ptr(), u.data_ptr(), stats.data_ptr(), B, L, C, KS, DIL, 1, st)
        dzn, dzw, dx = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)
        dWn, dWw = torch.empty(C, C, KS, device="cuda"), torch.empty(C, C, KS, device="cuda")
        dbn, dbw, dg2l = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda"), torch.zeros(B, C, device="cuda")
        ops._call("pb_conv_pair_bwd", du.data_ptr(), zn.data_ptr(), zw.data_ptr(), x.data_ptr(), wn_b.data_ptr(), ww_b.data_ptr(),
                  dzn.data_ptr(), dzw.data_ptr(), dx.data_ptr(), dWn.data_ptr(), dWw.data_ptr(), dbn.data_ptr(), dbw.data_ptr(),
                  dg2l.data_ptr(), B, L, C, KS, DIL, 1, *wsargs(pb), st)
        return zn, zw, u, stats, dx

    simt, tc = both_engines(pb, run, modes=(0, mode), op="conv")
    compare(["zn", "zw", "u", "stats", "dx"], simt, tc)


def test_token_head_tensor_core_path(pb):
    """bf16 token head through the tcgen05 GEMMs + batch softmax kernels vs an fp64 torch evaluation."""
    from proteinbert_b200 import ops
    g = torch.Generator().manual_seed(2)
    Bt, Lt, V = 6, 128, 26
    assert pb._lib.lib().pb_token_head_tc_ok(Bt, Lt, C, V)
    h = torch.randn(Bt, Lt, C, generator=g).bfloat16()
    Wo, bo = (torch.randn(V, C, generator=g) * 0.3), torch.randn(V, generator=g)
    hd, Wd, bd = h.cuda().requires_grad_(), Wo.cuda().requires_grad_(), bo.cuda().requires_grad_()
    p = ops.TokenHeadFn.apply(hd, Wd, bd)
    assert torch.allclose(p.sum(0), torch.ones(Lt, V, device="cuda"), atol=1e-5)
    hr = h.double().requires_grad_()
    Wr, br = Wo.bfloat16().double().requires_grad_(), bo.double().requires_grad_()
    pr = torch.softmax(hr @ Wr.t() + br, dim=0)
    assert rel_fro(p, pr) < 1e-3
    dP = torch.randn(Bt, Lt, V, generator=g)
    p.backward(dP.cuda())
    pr.backward(dP.double())
    assert rel_fro(hd.grad, hr.grad) < 1e-2
    assert rel_fro(Wd.grad, Wr.grad) < 1e-2
    assert float(bd.grad.abs().max()) < 1e-3 * float(Wd.grad.abs().max())     # mathematically zero


@pytest.mark.parametrize("M,N,K,act", [(64, 512, 8943, 1), (64, 8943, 512, 2), (5, 301, 64, 0), (100, 130, 70, 1), (128, 64, 200, 0)])
def test_annotation_layers_on_tensor_cores(pb, M, N, K, act):
    """pb_linear_fwd_tc / pb_linear_bwd_tc (tc_skinny.cu) against torch on the bf16-rounded operands (what the kernel
    multiplies) and against the fp32 entry points: odd leading dimensions, ragged tiles, every operand arrangement."""
    from proteinbert_b200 import ops
    lib = pb._lib.lib()
    assert lib.pb_linear_tc_ok(M, N, K)
    x = rnd(M, K, seed=1).cuda()
    if K > 1000:
        x = (x > 0.5).float()                      # annotations are 0/1
    W, b = rnd(N, K, scale=K ** -0.5, seed=2).cuda(), rnd(N, scale=0.1, seed=3).cuda()
    dy = rnd(M, N, scale=0.1, seed=4).cuda()
    st = torch.cuda.current_stream().cuda_stream
    ws = torch.empty(lib.pb_linear_tc_workspace(M, N, K), dtype=torch.uint8, device="cuda")
    z, y = torch.empty(M, N, device="cuda"), torch.empty(M, N, device="cuda")
    ops._call("pb_linear_fwd_tc", x.data_ptr(), W.data_ptr(), b.data_ptr(), z.data_ptr(), y.data_ptr(), M, N, K, act, ws.data_ptr(),
              ws.numel(), st)
    xr, Wr = x.bfloat16().double(), W.bfloat16().double()
    z_ref = xr @ Wr.t() + b.double()
    y_ref = {0: z_ref, 1: torch.nn.functional.gelu(z_ref), 2: torch.sigmoid(z_ref)}[act]
    assert rel_fro(z, z_ref) < 5e-5
    assert rel_fro(y, y_ref) < 5e-5
    assert rel_fro(z, x.double() @ W.double().t() + b.double()) < 1e-2          # vs the unrounded product
    zy = z if act == 1 else y
    dz, dx, dW, db = torch.empty(M, N, device="cuda"), torch.empty(M, K, device="cuda"), torch.empty(N, K, device="cuda"), torch.empty(N, device="cuda")
    ops._call("pb_linear_bwd_tc", dy.data_ptr(), zy.data_ptr() if act else None, x.data_ptr(), W.data_ptr(), dz.data_ptr(), dx.data_ptr(),
              dW.data_ptr(), db.data_ptr(), M, N, K, act, 0, ws.data_ptr(), ws.numel(), st)
    torch.cuda.synchronize()
    zd = z.double()
    dact = {0: torch.ones_like(zd), 1: None, 2: y.double() * (1 - y.double())}[act]
    if act == 1:
        zz = zd.clone().requires_grad_(True)
        torch.nn.functional.gelu(zz).backward(torch.ones_like(zz))
        dact = zz.grad
    dz_ref = dy.double() * dact
    assert rel_fro(dz, dz_ref) < 1e-5
    assert rel_fro(db, dz_ref.sum(0)) < 1e-5
    dzr = dz.bfloat16().double()
    assert rel_fro(dx, dzr @ Wr) < 5e-5
    assert rel_fro(dW, dzr.t() @ xr) < 5e-5
    assert rel_fro(dW, dz_ref.t() @ x.double()) < 1e-2
    # accumulate_dx adds into dX
    dx2 = dx.clone()
    ops._call("pb_linear_bwd_tc", dy.data_ptr(), zy.data_ptr() if act else None, x.data_ptr(), W.data_ptr(), dz.data_ptr(), dx2.data_ptr(),
              None, db.data_ptr(), M, N, K, act, 1, ws.data_ptr(), ws.numel(), st)
    assert rel_fro(dx2, 2 * dx.double()) < 1e-5
