"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle and the reference goldens.

Tolerances (stated once, used everywhere):
  fp32 mode : outputs / loss / gradients within 1e-4 relative (Frobenius) of the fp64 reference; an absolute
              floor for gradients that are mathematically zero (token-head bias, Wq/Wk) and 2e-3 for
              W_parameter, whose fp32 value depends on the association order.
  bf16 mode : loss <= 2e-3 rel, probabilities <= 1e-2 rel-Frobenius, concatenated gradient of the local-track
              tensors <= 5e-2 rel-Frobenius, each conv/linear/LayerNorm tensor <= 1e-1.
  integers  : bit-exact.
"""
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN, assert_close_rel, golden_batch, golden_params, load_model_golden, rel_fro
from oracle import data_oracle, model_oracle

pytestmark = pytest.mark.gpu

MATH_ZERO = ("pretraining_local_output.0.bias", "Wq_parameter", "Wk_parameter")
CANCELLATION_PRONE = ("W_parameter", "global_linear_layer", "global_norm", "global_to_local")


@pytest.fixture(scope="module")
def pb():
    import proteinbert_b200
    proteinbert_b200._lib.check(proteinbert_b200._lib.lib().pb_device_ok(0), "pb_device_ok")
    return proteinbert_b200


def build_model(pb, cfg, params, precision, dev="cuda"):
    model = pb.ProteinBERT(device=dev, precision=precision, **cfg)
    with torch.no_grad():
        for k, v in model.state_dict().items():
            v.copy_(params[k])
        for k, t in model.head_tensors().items():
            t.copy_(params[k])
    return model


def run_step(pb, model, batch, dev="cuda"):
    xl, xg, yl, yg, wl, wg = [t.to(dev) for t in batch]
    for t in model.head_tensors().values():
        t.grad = None
    for p in model.parameters():
        p.grad = None
    p_local, p_global = model({"local": xl, "global": xg})
    loss = pb.pretraining_loss(p_local, p_global, {"local": yl, "global": yg}, {"local": wl, "global": wg})
    loss.backward()
    grads = {k: p.grad for k, p in model.named_parameters()}
    # Wq / Wk cancel out of the reference layer exactly: the reference's autograd leaves
    # round-off noise in their .grad, this implementation leaves them untouched (None) -> compare as zeros
    grads.update({k: (t.grad if t.grad is not None else torch.zeros_like(t)) if ("Wq_" in k or "Wk_" in k) else t.grad
                  for k, t in model.head_tensors().items()})
    torch.cuda.synchronize()
    return p_local.detach(), p_global.detach(), loss.detach(), grads


def check_fp32(grads, ref_grads, what=""):
    gmax = max(float(v.abs().max()) for v in ref_grads.values())
    for k, ref in ref_grads.items():
        assert grads[k] is not None, k
        # W_parameter's gradient is a cancelling sum ~1e-4 of the largest gradient: fp32 atomics reorder it run to run,
        # so it gets the same absolute allowance as the mathematically-zero tensors on top of its relative bound
        floor = 1e-6 * gmax * np.sqrt(ref.numel()) if any(s in k for s in MATH_ZERO) or k.endswith("W_parameter") else 0.0
        rtol = 2e-3 if k.endswith("W_parameter") else 1e-4
        assert_close_rel(grads[k], ref, rtol, floor=floor, what=f"{what}{k}")


def test_golden_fp32_full_step(pb):
    gold = load_model_golden()
    cfg = {k: int(v) for k, v in gold["cfg"].items()}
    model = build_model(pb, cfg, golden_params(gold), "fp32")
    p_local, p_global, loss, grads = run_step(pb, model, golden_batch(gold))
    assert loss.dtype == torch.float64
    assert_close_rel(p_local, gold["out64"]["p_local"], 1e-4, what="p_local")
    assert_close_rel(p_global, gold["out64"]["p_global"], 1e-4, what="p_global")
    assert abs(float(loss) - float(gold["out64"]["loss"])) <= 1e-4 * abs(float(gold["out64"]["loss"]))
    check_fp32(grads, {k: torch.from_numpy(v) for k, v in gold["g64"].items()})


def test_golden_fp32_fused_global_track(pb):
    """Same step with the four-launch fused global-track kernels (ops_global_fused.cu) switched on."""
    gold = load_model_golden()
    cfg = {k: int(v) for k, v in gold["cfg"].items()}
    model = build_model(pb, cfg, golden_params(gold), "fp32")
    for blk in model.proteinBERT_blocks:
        blk.fuse_global = True
    p_local, p_global, loss, grads = run_step(pb, model, golden_batch(gold))
    assert_close_rel(p_global, gold["out64"]["p_global"], 1e-4, what="p_global")
    assert abs(float(loss) - float(gold["out64"]["loss"])) <= 1e-4 * abs(float(gold["out64"]["loss"]))
    check_fp32(grads, {k: torch.from_numpy(v) for k, v in gold["g64"].items()})


def test_golden_fp32_through_reference_style_loss(pb):
    """The reference loop applies torch's CrossEntropyLoss / BCELoss to the model outputs (utils.py:293-294);
    the drop-in must be differentiable through that route too."""
    gold = load_model_golden()
    cfg = {k: int(v) for k, v in gold["cfg"].items()}
    model = build_model(pb, cfg, golden_params(gold), "fp32")
    xl, xg, yl, yg, wl, wg = [t.cuda() for t in golden_batch(gold)]
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        pl, pg = model({"local": xl, "global": xg})
    ce, bce = torch.nn.CrossEntropyLoss(reduction="none"), torch.nn.BCELoss(reduction="none")
    loss = torch.mean(ce(pl.permute(0, 2, 1), yl) * wl) + torch.mean(bce(pg, yg.float()) * wg)
    loss.backward()
    assert abs(float(loss) - float(gold["out64"]["loss"])) <= 1e-4 * abs(float(gold["out64"]["loss"]))
    k = "proteinBERT_blocks.0.local_narrow_conv_layer.0.weight"
    assert_close_rel(dict(model.named_parameters())[k].grad, gold["g64"][k], 1e-4, what=k)
    k = "global_linear_layer.0.weight"
    assert_close_rel(dict(model.named_parameters())[k].grad, gold["g64"][k], 1e-4, what=k)


def check_bf16(p_local, p_global, loss, grads, ref):
    ref_pl, ref_pg, ref_loss, ref_grads = ref
    assert abs(float(loss) - float(ref_loss)) <= 2e-3 * abs(float(ref_loss))
    assert rel_fro(p_local, ref_pl) <= 1e-2
    assert rel_fro(p_global, ref_pg) <= 1e-2
    local = [k for k in ref_grads if not any(s in k for s in CANCELLATION_PRONE + MATH_ZERO) and "heads" not in k]
    cat = torch.cat([grads[k].double().cpu().reshape(-1) for k in local])
    cat_ref = torch.cat([ref_grads[k].double().reshape(-1) for k in local])
    assert rel_fro(cat, cat_ref) <= 5e-2
    for k in local:
        assert rel_fro(grads[k], ref_grads[k]) <= 1e-1, (k, rel_fro(grads[k], ref_grads[k]))
    for k, g in grads.items():
        assert g is not None and torch.isfinite(g).all(), k


def test_golden_bf16_full_step(pb):
    gold = load_model_golden()
    cfg = {k: int(v) for k, v in gold["cfg"].items()}
    model = build_model(pb, cfg, golden_params(gold), "bf16")
    out = run_step(pb, model, golden_batch(gold))
    ref = (gold["out64"]["p_local"], gold["out64"]["p_global"], gold["out64"]["loss"],
           {k: torch.from_numpy(v) for k, v in gold["g64"].items()})
    check_bf16(*out, ref)


# the width the tensor-core tiling is built for (C=128, G=512), short enough for the CPU oracle
WIDE = dict(sequences_length=256, num_annotations=517, local_dim=128, global_dim=512, key_dim=64, num_heads=4, num_blocks=2)


@pytest.fixture(scope="module")
def wide_ref():
    params = model_oracle.init_params(seed=11, **WIDE)
    g = torch.Generator().manual_seed(5)
    for k in params:
        if "norm" in k:
            params[k] = params[k] + (0.2 if k.endswith("weight") else 0.1) * torch.randn(params[k].shape, generator=g)
    batch = model_oracle.synthetic_batch(6, WIDE["sequences_length"], WIDE["num_annotations"], seed=3)
    p64 = {k: v.double() for k, v in params.items()}
    b64 = (batch[0], batch[1].double(), batch[2], batch[3].double(), batch[4], batch[5])
    return params, batch, model_oracle.loss_and_grads(p64, b64, attention="closed_form")


@pytest.mark.parametrize("fuse_global", [False, True])
def test_wide_fp32(pb, wide_ref, fuse_global):
    params, batch, ref = wide_ref
    model = build_model(pb, WIDE, params, "fp32")
    for blk in model.proteinBERT_blocks:
        blk.fuse_global = fuse_global
    p_local, p_global, loss, grads = run_step(pb, model, batch)
    assert_close_rel(p_local, ref[0], 1e-4, what="p_local")
    assert_close_rel(p_global, ref[1], 1e-4, what="p_global")
    assert abs(float(loss) - float(ref[2])) <= 1e-4 * abs(float(ref[2]))
    check_fp32(grads, ref[3], "wide/")


def test_wide_bf16_tensor_core_path(pb, wide_ref):
    params, batch, ref = wide_ref
    model = build_model(pb, WIDE, params, "bf16")
    out = run_step(pb, model, batch)
    check_bf16(*out, ref)


def test_reference_module_surface(pb):
    """dict-in/dict-out blocks, channels-first I/O, standalone attention layer vs the as-executed oracle."""
    gold = load_model_golden()
    cfg = {k: int(v) for k, v in gold["cfg"].items()}
    params = golden_params(gold, torch.float64)
    model = build_model(pb, cfg, golden_params(gold), "fp32")
    B, L, C, G = 4, cfg["sequences_length"], cfg["local_dim"], cfg["global_dim"]
    g = torch.Generator().manual_seed(0)
    x = torch.randn(B, C, L, generator=g)
    gl = torch.randn(B, G, generator=g)
    blk = model.proteinBERT_blocks[1]
    out = blk({"local": x.cuda(), "global": gl.cuda()})
    assert out["local"].shape == (B, C, L) and out["global"].shape == (B, G)
    h_ref, g_ref = model_oracle.block_forward(params, 1, x.double().permute(0, 2, 1), gl.double())
    assert_close_rel(out["local"].permute(0, 2, 1), h_ref, 1e-4, what="block local")
    assert_close_rel(out["global"], g_ref, 1e-4, what="block global")
    att = blk.global_attention_layer
    hl = torch.randn(B, L, C, generator=g)
    a = att({"local": hl.cuda(), "global": gl.cuda()})
    a_ref = model_oracle.attention_as_executed(hl.double(), gl.double(), model_oracle.heads_of(params, 1),
                                               params["proteinBERT_blocks.1.global_attention_layer.W_parameter"])
    assert_close_rel(a, a_ref, 1e-4, what="GlobalAttention")
    head = att.global_attention_heads[2]
    o = head({"local": hl.cuda(), "global": gl.cuda()})
    assert o.shape == (B, cfg["key_dim"], G // cfg["num_heads"])
    wq, wk, wv = model_oracle.heads_of(params, 1)[2]
    # as executed by the reference head (modules.py:49-60): softmax over the repeated-query axis, then P @ V
    K = cfg["key_dim"]
    q = torch.tanh(gl.double().unsqueeze(1).expand(-1, K, -1) @ wq)
    s = q @ torch.tanh(hl.double() @ wk).transpose(1, 2) / K ** 0.5
    o_ref = torch.softmax(s, dim=1) @ model_oracle.gelu(hl.double() @ wv)
    assert_close_rel(o, o_ref, 1e-4, what="GlobalAttentionHead")


def test_conv_pair_against_torch_conv1d(pb):
    """The kernel in isolation against nn.functional.conv1d (fp32, odd sizes, dilation 5 and 3)."""
    from proteinbert_b200 import ops
    lib, check = pb._lib.lib(), pb._lib.check
    g = torch.Generator().manual_seed(1)
    for (B, L, C, ks, dil) in [(3, 50, 24, 9, 5), (2, 130, 128, 9, 5), (1, 17, 8, 5, 3)]:
        x = torch.randn(B, L, C, generator=g)
        wn = torch.randn(C, C, ks, generator=g) * 0.1
        ww = torch.randn(C, C, ks, generator=g) * 0.1
        bn, bw = torch.randn(C, generator=g), torch.randn(C, generator=g)
        g2l = torch.randn(B, C, generator=g)
        xd, bnd, bwd, g2ld = x.cuda(), bn.cuda(), bw.cuda(), g2l.cuda()      # keep the device tensors alive
        wn_f, _ = ops.conv_pack(wn.cuda(), 0)
        ww_f, _ = ops.conv_pack(ww.cuda(), 0)
        zn, zw, u = torch.empty_like(xd), torch.empty_like(xd), torch.empty_like(xd)
        stats = torch.zeros(B, 2, dtype=torch.float64, device="cuda")
        check(lib.pb_conv_pair_fwd(xd.data_ptr(), wn_f.data_ptr(), ww_f.data_ptr(), bnd.data_ptr(), bwd.data_ptr(),
                                   g2ld.data_ptr(), zn.data_ptr(), zw.data_ptr(), u.data_ptr(), stats.data_ptr(),
                                   B, L, C, ks, dil, 0, torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        xc = x.double().permute(0, 2, 1)
        rn = torch.nn.functional.conv1d(xc, wn.double(), bn.double(), padding="same").permute(0, 2, 1)
        rw = torch.nn.functional.conv1d(xc, ww.double(), bw.double(), padding="same", dilation=dil).permute(0, 2, 1)
        ru = x.double() + model_oracle.gelu(rn) + model_oracle.gelu(rw) + g2l.double().unsqueeze(1)
        assert_close_rel(zn, rn, 1e-5, what="narrow")
        assert_close_rel(zw, rw, 1e-5, what="wide")
        assert_close_rel(u, ru, 1e-5, what="u")
        assert_close_rel(stats[:, 0], ru.sum((1, 2)), 1e-5, floor=1e-3)
        assert_close_rel(stats[:, 1], (ru * ru).sum((1, 2)), 1e-5)


@pytest.mark.parametrize("dtype,two,shape", [("fp32", True, (5, 96, 32)), ("fp32", False, (3, 40, 64)),
                                            ("bf16", True, (19, 256, 128)), ("bf16", False, (6, 128, 256)),
                                            # L*C = 65536: the one-launch cluster kernel, two vectors per thread; 150 samples =
                                            # several rounds per cluster
                                            ("bf16", True, (9, 512, 128)), ("bf16", False, (150, 512, 128)),
                                            ("bf16", True, (64, 512, 128))])
def test_fused_layernorm_gelu_backward(pb, dtype, two, shape):
    """pb_ln_joint_bwd_gelu against torch autograd (double) of u -> LayerNorm((L,C)) with z -> gelu side branches
    (modules.py:207-217 in reverse): du, dz, dgamma/dbeta, channel sums and the per-sample sums."""
    from proteinbert_b200 import ops
    B, L, C = shape
    assert pb._lib.lib().pb_ln_joint_bwd_gelu_ok(B, L, C)
    g = torch.Generator().manual_seed(B * 1000 + L)
    act = torch.float32 if dtype == "fp32" else torch.bfloat16
    dt = 0 if dtype == "fp32" else 1
    u, dy, z0, z1 = (torch.randn(B, L, C, generator=g).cuda().to(act) for _ in range(4))
    gamma = (1 + 0.2 * torch.randn(L, C, generator=g)).cuda()
    stats = torch.stack([u.double().sum((1, 2)), (u.double() ** 2).sum((1, 2))], 1).contiguous()
    du, dz0, dz1 = torch.empty_like(u), torch.empty_like(u), torch.empty_like(u)
    dgamma, dbeta = torch.zeros(L, C, device="cuda"), torch.zeros(L, C, device="cuda")
    dsample, db0, db1 = torch.zeros(B, C, device="cuda"), torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    sums = torch.zeros(B, 2, dtype=torch.float64, device="cuda")
    ops._call("pb_ln_joint_bwd_gelu", dy.data_ptr(), u.data_ptr(), stats.data_ptr(), gamma.data_ptr(), z0.data_ptr(),
              z1.data_ptr() if two else None, du.data_ptr(), dz0.data_ptr(), dz1.data_ptr() if two else None, dgamma.data_ptr(),
              dbeta.data_ptr(), dsample.data_ptr(), db0.data_ptr(), db1.data_ptr() if two else None, sums.data_ptr(), B, L, C,
              1e-5, dt, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ud = u.double().requires_grad_(True)
    gd = gamma.double().requires_grad_(True)
    bd = torch.zeros(L, C, dtype=torch.float64, device="cuda", requires_grad=True)
    y = torch.nn.functional.layer_norm(ud, (L, C), gd, bd, 1e-5)
    y.backward(dy.double())
    du_ref = ud.grad
    tol = 1e-5 if dtype == "fp32" else 4e-3
    assert_close_rel(du, du_ref, tol, what="du")
    assert_close_rel(dgamma, gd.grad, tol, what="dgamma")
    assert_close_rel(dbeta, bd.grad, tol, what="dbeta")
    # the GELU branches act on the STORED du (rounded to the activation type), as the unfused kernels do
    dus = du.double()
    for z, dz, db in ((z0, dz0, db0), (z1, dz1, db1)) if two else ((z0, dz0, db0),):
        zd = z.double().requires_grad_(True)
        torch.nn.functional.gelu(zd).backward(dus)
        assert_close_rel(dz, zd.grad, tol, what="dz")
        assert_close_rel(db, dz.double().sum((0, 1)), 1e-5 if dtype == "fp32" else 1e-4, floor=1e-4, what="db")
    assert_close_rel(dsample, dus.sum(1), 1e-5 if dtype == "fp32" else 1e-4, floor=1e-4, what="dsample")


def test_token_head_batch_axis_softmax(pb):
    g = torch.Generator().manual_seed(2)
    B, L, C, V = 5, 33, 48, 26
    from proteinbert_b200 import ops
    h = torch.randn(B, L, C, generator=g)
    Wo, bo = torch.randn(V, C, generator=g) * 0.3, torch.randn(V, generator=g)
    hd, Wd, bd = h.cuda().requires_grad_(), Wo.cuda().requires_grad_(), bo.cuda().requires_grad_()
    p = ops.TokenHeadFn.apply(hd, Wd, bd)
    assert torch.allclose(p.sum(0), torch.ones(L, V, device="cuda"), atol=1e-5)        # normalised over the batch
    hr, Wr, br = h.double().requires_grad_(), Wo.double().requires_grad_(), bo.double().requires_grad_()
    pr = torch.softmax(hr @ Wr.t() + br, dim=0)
    assert_close_rel(p, pr, 1e-5)
    dP = torch.randn(B, L, V, generator=g)
    p.backward(dP.cuda())
    pr.backward(dP.double())
    assert_close_rel(hd.grad, hr.grad, 1e-4)
    assert_close_rel(Wd.grad, Wr.grad, 1e-4)
    assert float(bd.grad.abs().max()) < 1e-5                                            # mathematically zero
    one = ops.TokenHeadFn.apply(torch.randn(1, 4, C, device="cuda"), Wd.detach(), bd.detach())
    assert torch.equal(one, torch.ones_like(one))                                       # B = 1 -> all ones


def test_loss_matches_torch_losses_including_clamp(pb):
    g = torch.Generator().manual_seed(4)
    B, L, V, A = 3, 21, 26, 77
    pl = torch.softmax(torch.randn(B, L, V, generator=g), 0)
    pg = torch.sigmoid(torch.randn(B, A, generator=g) * 4)
    pg[0, 0], pg[0, 1], pg[1, 2] = 0.0, 1.0, 1.0          # exercises the -100 clamp and the 1e-12 floor
    yl = torch.randint(0, V, (B, L), generator=g)
    yg = (torch.rand(B, A, generator=g) < 0.3).float()
    wl, wg = torch.rand(B, L, generator=g).double(), torch.rand(B, A, generator=g).double()
    pld, pgd = pl.cuda().requires_grad_(), pg.cuda().requires_grad_()
    loss = pb.pretraining_loss(pld, pgd, {"local": yl.cuda(), "global": yg.cuda()}, {"local": wl.cuda(), "global": wg.cuda()})
    loss.backward()
    plr, pgr = pl.clone().requires_grad_(), pg.clone().requires_grad_()
    ce, bce = torch.nn.CrossEntropyLoss(reduction="none"), torch.nn.BCELoss(reduction="none")
    ref = torch.mean(ce(plr.permute(0, 2, 1), yl) * wl) + torch.mean(bce(pgr, yg) * wg)
    ref.backward()
    assert loss.dtype == torch.float64
    assert abs(float(loss) - float(ref)) <= 1e-6 * abs(float(ref))
    assert_close_rel(pld.grad, plr.grad, 1e-5)
    assert_close_rel(pgd.grad, pgr.grad, 1e-5)


def test_embedding_forward_backward(pb):
    from proteinbert_b200 import ops
    g = torch.Generator().manual_seed(6)
    tok = torch.randint(0, 26, (7, 40), generator=g)
    tok[:, -5:] = 0
    E = torch.randn(26, 64, generator=g)
    Ed = E.cuda().requires_grad_()
    for dt, adt in ((0, torch.float32), (1, torch.bfloat16)):
        Ed.grad = None
        out = ops.EmbeddingFn.apply(tok.cuda(), Ed, dt)
        assert out.dtype == adt and torch.equal(out.float().cpu(), E[tok].to(adt).float())
        d = torch.randn(7, 40, 64, generator=g).to(adt)
        out.backward(d.cuda())
        ref = torch.zeros(26, 64, dtype=torch.float64).index_add_(0, tok.reshape(-1), d.double().reshape(-1, 64))
        assert_close_rel(Ed.grad, ref, 1e-5)


def test_corrupt_pad_bit_exact_against_reference_records(pb):
    """Every golden record (real reference run) through the CUDA kernel, as one ragged batch."""
    z = np.load(os.path.join(GOLDEN, "data_path.npz"))
    L, A, n = int(z["L"]), int(z["A"]), int(z["n"])
    seqs = [z[f"s{r}/seq"].tobytes().decode("latin-1") for r in range(n)]
    anns = [z[f"s{r}/ann"] for r in range(n)]
    draws = []
    for r in range(n):
        pre = f"s{r}/"
        noised = float(z[pre + "coin"]) > 0.5
        draws.append({"crop_start": int(z[pre + "crop"]), "tok_mask": z[pre + "tok_mask"], "tok_rand": z[pre + "tok_rand"],
                      "coin": float(z[pre + "coin"]), "keep": z[pre + "keep"] if noised else None,
                      "add": z[pre + "add"] if noised else None})
    X, Y, W = pb.data_processing.corrupt_pad(pb.data_processing.pack_batch(seqs, anns, draws, L))
    torch.cuda.synchronize()
    for name, got in (("x_local", X["local"]), ("x_global", X["global"]), ("y_local", Y["local"]),
                      ("y_global", Y["global"]), ("w_local", W["local"]), ("w_global", W["global"])):
        got = got.cpu().numpy()
        for r in range(n):
            ref = z[f"s{r}/{name}"]
            assert got[r].dtype == ref.dtype and np.array_equal(got[r], ref), (name, r)


def test_corrupt_pad_against_oracle_at_full_size(pb):
    """B=64, L=512, A=8943 with fresh draws: CUDA vs the numpy oracle, bit-exact, plus invariants."""
    import random
    rnd = random.Random(7777)
    B, L, A = 64, 512, 8943
    letters = data_oracle.ALL_AMINO_ACIDS
    seqs = ["".join(letters[rnd.randint(0, 21)] for _ in range(rnd.randint(0, 2 * L))) for _ in range(B)]
    anns = (np.random.RandomState(0).rand(B, A) < 0.005).astype(np.int64)
    anns[3] = 0
    torch.manual_seed(0)
    draws = [pb.data_processing.draw_sample(len(s), A, L) for s in seqs]
    packed = pb.data_processing.pack_batch(seqs, anns, draws, L)
    X, Y, W = pb.data_processing.corrupt_pad(packed)
    # out=: the same six tensors, overwritten in place (poisoned first)
    for t in (X["local"], X["global"], Y["local"], Y["global"], W["local"], W["global"]):
        t.fill_(7)
    X2, Y2, W2 = pb.data_processing.corrupt_pad(packed, out=(X, Y, W))
    assert X2["local"].data_ptr() == X["local"].data_ptr() and W2["global"].data_ptr() == W["global"].data_ptr()
    nd = [dict(crop_start=d["crop_start"], tok_mask=d["tok_mask"].numpy(), tok_rand=d["tok_rand"].numpy(), coin=d["coin"],
               keep=(d["keep"].numpy() if d["keep"] is not None else np.zeros(A, np.int64)),
               add=(d["add"].numpy() if d["add"] is not None else np.zeros(A, np.int64))) for d in draws]
    ref = data_oracle.corrupt_batch(seqs, anns, L, nd)
    for got, r in zip((X["local"], X["global"], Y["local"], Y["global"], W["local"], W["global"]), ref):
        assert np.array_equal(got.cpu().numpy(), r)
    yl = Y["local"].cpu()
    assert ((yl == 0) == (W["local"].cpu() == 0)).all()
    assert float(W["global"][3].sum()) == 0.0 and set(np.unique(X["global"].cpu().numpy())) <= {0.0, 1.0, 2.0}


def test_full_size_properties_bf16(pb):
    """Config-2 sizes (B=64, L=512, 128/512, 6 blocks, 8943 annotations), bf16: properties that do not need
    an oracle run -- batch-axis normalisation, finite gradients for all 133 tensors, exact zeros where the
    data says so, and agreement of two identical steps."""
    cfg = dict(sequences_length=512, num_annotations=8943, local_dim=128, global_dim=512, key_dim=64, num_heads=4, num_blocks=6)
    torch.manual_seed(0)
    model = pb.ProteinBERT(device="cuda", precision="bf16", **cfg)
    batch = model_oracle.synthetic_batch(64, 512, 8943, seed=1)
    p_local, p_global, loss, grads = run_step(pb, model, batch)
    assert torch.allclose(p_local.sum(0), torch.ones_like(p_local[0]), atol=1e-4)
    assert float(p_global.min()) >= 0.0 and float(p_global.max()) <= 1.0 and torch.isfinite(loss)
    named = dict(model.named_parameters())
    assert len(named) == 133
    for k in named:
        assert grads[k] is not None and torch.isfinite(grads[k]).all(), k
        assert float(grads[k].abs().max()) > 0 or k == "pretraining_local_output.0.bias", k
    used = torch.bincount(batch[0].reshape(-1), minlength=26) > 0
    ge = grads["local_embedding.weight"].cpu()
    assert (ge[~used] == 0).all() and (ge[used].abs().sum(1) > 0).all()
    _, _, loss2, grads2 = run_step(pb, model, batch)
    assert abs(float(loss2) - float(loss)) <= 1e-6 * abs(float(loss))
    k = "proteinBERT_blocks.3.local_wide_conv_layer.0.weight"
    assert rel_fro(grads2[k], grads[k]) < 1e-3


def test_graphed_step_and_two_stage_backward_match_eager(pb, wide_ref):
    """GraphedPretrainStep (one CUDA graph) and its data-parallel form (backward split into two graphs with a flat
    gradient bucket per half; exercised here on one GPU) must leave the same gradients as the eager step."""
    params, batch, _ = wide_ref
    model = build_model(pb, WIDE, params, "bf16")
    xl, xg, yl, yg, wl, wg = [t.cuda() for t in batch]
    X, Y, W = {"local": xl, "global": xg}, {"local": yl, "global": yg}, {"local": wl, "global": wg}
    loss_e = pb.pretraining_step(model, X, Y, W)
    ref = {n: p.grad.clone() for n, p in model.named_parameters()}
    for kwargs in ({}, {"force_split": True}, {"force_bucket": True}):
        step = pb.GraphedPretrainStep(model, X, Y, W, warmup=1, **kwargs)
        if kwargs:        # every gradient was written by its kernel straight into the flat bucket (no copies needed)
            assert step.bucket is not None and step.bucket.loose == [], step.bucket.loose
            for n, p in model.named_parameters():
                assert p.grad.data_ptr() == step.bucket.view(n).data_ptr(), n

        loss_g = step(X, Y, W)
        torch.cuda.synchronize()
        assert abs(float(loss_g) - float(loss_e)) <= 1e-6 * abs(float(loss_e))
        for n, p in model.named_parameters():
            assert rel_fro(p.grad, ref[n]) < 2e-3 or float(ref[n].abs().max()) < 1e-6, (kwargs, n, rel_fro(p.grad, ref[n]))
        if kwargs:
            assert step.ddp_check(X, Y, W) == 0.0          # one rank: the exchange is the identity
        # host batches with one step of look-ahead: the staged copy of the next batch must be the one the step consumes
        hostA = [t.cpu().pin_memory() for t in batch]
        hostB = [t.cpu().pin_memory() for t in batch]
        hostB[1] = (1.0 - hostB[1]).pin_memory()                    # a different annotation input -> a different loss
        dA = ({"local": hostA[0], "global": hostA[1]}, {"local": hostA[2], "global": hostA[3]}, {"local": hostA[4], "global": hostA[5]})
        dB = ({"local": hostB[0], "global": hostB[1]}, {"local": hostB[2], "global": hostB[3]}, {"local": hostB[4], "global": hostB[5]})
        lA = float(step(*dA, prefetch=dB))
        lB = float(step(*dB, prefetch=dA))
        lA2 = float(step(*dA))
        lB_direct = float(step(*dB))
        tol = 1e-6 * abs(float(loss_e))                           # float atomics: replays agree to rounding, not bitwise
        assert abs(lA - float(loss_e)) <= tol and abs(lA2 - lA) <= tol
        assert abs(lB - lB_direct) <= tol and abs(lB - lA) > 100 * tol
        # losses read one step late through pinned memory: handle k holds step k's loss even after step k+1 was issued
        step(*dA)
        hA = step.read_loss_async()
        step(*dB)
        hB = step.read_loss_async()
        assert abs(hA.value() - lA) <= tol and abs(hB.value() - lB) <= tol
        del step


@pytest.mark.parametrize("name,B,cfg", [
    ("C4: L=2048", 4, dict(sequences_length=2048, num_annotations=8943, local_dim=128, global_dim=512, key_dim=64, num_heads=4, num_blocks=2)),
    ("C5: C=256 G=1024 L=1024", 2, dict(sequences_length=1024, num_annotations=8943, local_dim=256, global_dim=1024, key_dim=64, num_heads=4, num_blocks=2)),
])
def test_other_baseline_configs_run_in_both_precisions(pb, name, B, cfg):
    """BASELINE.json's long-sequence and wide configurations (block count cut to 2 for test time): the bf16 path (tensor
    cores where the tiling covers the shape, the SIMT engine elsewhere) must agree with the fp32 path on the loss and
    leave finite gradients."""
    losses = {}
    for prec in ("fp32", "bf16"):
        torch.manual_seed(0)
        model = pb.ProteinBERT(device=torch.device("cuda"), precision=prec, **cfg)
        b = [t.cuda() for t in model_oracle.synthetic_batch(B, cfg["sequences_length"], cfg["num_annotations"], seed=1)]
        X, Y, W = {"local": b[0], "global": b[1]}, {"local": b[2], "global": b[3]}, {"local": b[4], "global": b[5]}
        losses[prec] = float(pb.pretraining_step(model, X, Y, W))
        assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters()), (name, prec)
    assert abs(losses["bf16"] - losses["fp32"]) <= 2e-3 * abs(losses["fp32"]), (name, losses)


def test_fused_adam_matches_torch_adam(pb):
    """pb_adam_step (FusedAdam) against torch.optim.Adam on the same parameters and gradients for several steps, odd sizes
    and a tensor above the 64k-element piece size; interchangeable state_dict."""
    g = torch.Generator().manual_seed(0)
    shapes = [(7,), (33, 5), (300, 301), (128, 128, 9), (1,)]
    p_ref = [torch.randn(s, generator=g).cuda().requires_grad_(True) for s in shapes]
    p_new = [p.detach().clone().requires_grad_(True) for p in p_ref]
    ref = torch.optim.Adam(p_ref, lr=2e-4, foreach=False, fused=False)
    new = pb.FusedAdam(p_new, lr=2e-4)
    for step in range(4):
        for a, b_ in zip(p_ref, p_new):
            gr = torch.randn(a.shape, generator=g).cuda() * (10.0 ** (step - 2))
            a.grad, b_.grad = gr.clone(), gr.clone()
        ref.step(); new.step()
        for a, b_ in zip(p_ref, p_new):
            assert_close_rel(b_.detach(), a.detach(), 1e-6, what=f"param step {step}")
    for a, b_ in zip(p_ref, p_new):
        assert_close_rel(new.state[b_]["exp_avg"], ref.state[a]["exp_avg"], 1e-6)
        assert_close_rel(new.state[b_]["exp_avg_sq"], ref.state[a]["exp_avg_sq"], 1e-6)
    ref2 = torch.optim.Adam(p_ref, lr=2e-4)
    ref2.load_state_dict(new.state_dict())                      # same keys / layout as the torch optimizer
    for a in p_ref:
        assert float(ref2.state[a]["step"]) == float(ref.state[a]["step"]) == 4


def test_fused_adam_survives_load_state_dict_after_stepping(pb):
    """step, save, step, load, step: torch replaces the moment tensors in load_state_dict, so the cached device table of
    raw pointers must be rebuilt and the loaded step count honoured -- compared with torch.optim.Adam doing the same."""
    g = torch.Generator().manual_seed(1)
    shapes = [(9,), (70, 130), (128, 128, 9)]
    p_ref = [torch.randn(s, generator=g).cuda().requires_grad_(True) for s in shapes]
    p_new = [p.detach().clone().requires_grad_(True) for p in p_ref]
    ref = torch.optim.Adam(p_ref, lr=1e-3, foreach=False, fused=False)
    new = pb.FusedAdam(p_new, lr=1e-3)
    grads = [[torch.randn(s, generator=g).cuda() for s in shapes] for _ in range(3)]

    def give(i):
        for a, b_, gr in zip(p_ref, p_new, grads[i]):      # the SAME gradient tensors every time: their addresses stay put
            if a.grad is None:
                a.grad, b_.grad = gr.clone(), gr.clone()
            else:
                a.grad.copy_(gr); b_.grad.copy_(gr)

    give(0); ref.step(); new.step()
    saved_ref, saved_new = ref.state_dict(), new.state_dict()
    import copy
    saved_ref, saved_new = copy.deepcopy(saved_ref), copy.deepcopy(saved_new)
    saved_p = [p.detach().clone() for p in p_ref]
    give(1); ref.step(); new.step()
    with torch.no_grad():                                   # back to the checkpoint: weights and optimizer state
        for a, b_, s in zip(p_ref, p_new, saved_p):
            a.copy_(s); b_.copy_(s)
    ref.load_state_dict(saved_ref); new.load_state_dict(saved_new)
    give(2); ref.step(); new.step()
    for a, b_ in zip(p_ref, p_new):
        assert_close_rel(b_.detach(), a.detach(), 1e-6, what="param after load")
        assert_close_rel(new.state[b_]["exp_avg"], ref.state[a]["exp_avg"], 1e-6)
        assert_close_rel(new.state[b_]["exp_avg_sq"], ref.state[a]["exp_avg_sq"], 1e-6)
    assert all(float(v["step"]) == 2 for v in new.state_dict()["state"].values())
