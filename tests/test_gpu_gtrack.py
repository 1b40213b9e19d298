"""GPU: the cluster-unit global-track kernels (csrc/gtrack.cu, ops.GTrackFn) against torch autograd in float64 of the
reference formulas (modules.py:92,208,221-229), forward values and every gradient, for a full unit [post of block i ; pre of
block i+1] and for the two edge units (pre only, post only).  Tolerance: fp32 parity, 1e-4 relative (the products are
error-compensated 3xTF32; measured ~1e-6)."""
import pytest
import torch

from helpers import assert_close_rel

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pb():
    import proteinbert_b200
    return proteinbert_b200


def gelu(t):
    return torch.nn.functional.gelu(t)


def ln(v, gamma, beta):
    return torch.nn.functional.layer_norm(v, (v.shape[-1],), gamma, beta, 1e-5)


def make(B, G, C, K, seed):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s, scale=1.0: (torch.randn(*s, generator=g) * scale)
    return dict(g=r(B, G), gh=r(B, G, scale=0.3), pool=r(B, G, scale=20.0), Wp=r(K), gamma1=1 + 0.2 * r(G), beta1=0.1 * r(G),
                W4=r(G, G, scale=G ** -0.5), b4=0.1 * r(G), gamma2=1 + 0.2 * r(G), beta2=0.1 * r(G),
                W1=r(C, G, scale=G ** -0.5), b1=0.1 * r(C), W3=r(G, G, scale=G ** -0.5), b3=0.1 * r(G),
                d_g2=r(B, G, scale=0.1), d_g2l=r(B, C, scale=0.1), d_gh=r(B, G, scale=0.1))


@pytest.mark.parametrize("B,G,C,K", [(64, 512, 128, 64), (5, 64, 32, 16), (16, 512, 128, 64), (33, 256, 64, 64)])
@pytest.mark.parametrize("mode", ["full", "pre", "post"])
def test_gtrack_unit_against_float64_autograd(pb, B, G, C, K, mode):
    from proteinbert_b200 import ops
    if not ops.gtrack_ok(B, G, C):
        pytest.skip("no 16-CTA cluster on this device / shape not covered")
    t = make(B, G, C, K, seed=B + G)
    has_post, has_pre = mode != "pre", mode != "post"
    post_names = ("gh", "pool", "Wp", "gamma1", "beta1", "W4", "b4", "gamma2", "beta2")
    pre_names = ("W1", "b1", "W3", "b3")
    dev = {k: v.cuda().requires_grad_(not k.startswith("d_")) for k, v in t.items()}
    ref = {k: v.double().requires_grad_(not k.startswith("d_")) for k, v in t.items()}
    args = [dev["g"]] + [dev[n] if has_post else None for n in post_names] + [dev[n] if has_pre else None for n in pre_names]
    g2, g2l, ghn = ops.GTrackFn.apply(*args, C, None, None)
    # float64 reference
    x = ref["g"]
    if has_post:
        coef = ref["Wp"].sum() / K
        g1 = ln(ref["g"] + ref["gh"] + coef * ref["pool"], ref["gamma1"], ref["beta1"])
        x = ln(g1 + gelu(g1 @ ref["W4"].t() + ref["b4"]), ref["gamma2"], ref["beta2"])
        assert_close_rel(g2, x, 1e-5, what="g2")
    outs, douts, routs = [], [], []
    if has_post:
        outs.append(g2); douts.append(dev["d_g2"]); routs.append((x, ref["d_g2"]))
    if has_pre:
        r_g2l = gelu(x @ ref["W1"].t() + ref["b1"])
        r_gh = gelu(x @ ref["W3"].t() + ref["b3"])
        assert_close_rel(g2l, r_g2l, 1e-5, what="g2l")
        assert_close_rel(ghn, r_gh, 1e-5, what="gh")
        outs += [g2l, ghn]; douts += [dev["d_g2l"], dev["d_gh"]]
        routs += [(r_g2l, ref["d_g2l"]), (r_gh, ref["d_gh"])]
    torch.autograd.backward(outs, douts)
    ops.join_side_streams()
    torch.cuda.synchronize()
    torch.autograd.backward([a for a, _ in routs], [b for _, b in routs])
    names = ["g"] + (list(post_names) if has_post else []) + (list(pre_names) if has_pre else [])
    for n in names:
        assert dev[n].grad is not None, n
        floor = 1e-6 * float(ref[n].grad.abs().max()) * ref[n].grad.numel() ** 0.5
        assert_close_rel(dev[n].grad, ref[n].grad, 1e-4, floor=floor, what=f"d{n}")


def test_chained_units_use_the_dz3_link(pb):
    """Two units in a row, as ProteinBERT.forward chains them: the second unit's backward hands dz3 to the first through
    the link; gradients must equal float64 autograd through both."""
    from proteinbert_b200 import ops
    B, G, C, K = 64, 512, 128, 64
    if not ops.gtrack_ok(B, G, C):
        pytest.skip("no 16-CTA cluster on this device")
    a, b = make(B, G, C, K, seed=1), make(B, G, C, K, seed=2)
    post_names = ("Wp", "gamma1", "beta1", "W4", "b4", "gamma2", "beta2")
    pre_names = ("W1", "b1", "W3", "b3")
    da = {k: v.cuda().requires_grad_(not k.startswith("d_")) for k, v in a.items()}
    db = {k: v.cuda().requires_grad_(not k.startswith("d_")) for k, v in b.items()}
    ra = {k: v.double().requires_grad_(not k.startswith("d_")) for k, v in a.items()}
    rb = {k: v.double().requires_grad_(not k.startswith("d_")) for k, v in b.items()}
    links = [ops.GTrackLink() for _ in range(3)]
    g2, g2l, gh = ops.GTrackFn.apply(da["g"], da["gh"], da["pool"], *[da[n] for n in post_names], *[da[n] for n in pre_names], C, links[0], links[1])
    g3, _, _ = ops.GTrackFn.apply(g2, gh, db["pool"], *[db[n] for n in post_names], None, None, None, None, C, links[1], links[2])
    torch.autograd.backward([g3, g2l], [db["d_g2"], da["d_g2l"]])
    ops.join_side_streams()
    torch.cuda.synchronize()

    def post(r, g, gh, pool):
        g1 = ln(g + gh + (r["Wp"].sum() / K) * pool, r["gamma1"], r["beta1"])
        return ln(g1 + gelu(g1 @ r["W4"].t() + r["b4"]), r["gamma2"], r["beta2"])

    x2 = post(ra, ra["g"], ra["gh"], ra["pool"])
    r_g2l, r_gh = gelu(x2 @ ra["W1"].t() + ra["b1"]), gelu(x2 @ ra["W3"].t() + ra["b3"])
    x3 = post(rb, x2, r_gh, rb["pool"])
    assert_close_rel(g3, x3, 1e-5, what="g3")
    torch.autograd.backward([x3, r_g2l], [rb["d_g2"], ra["d_g2l"]])
    for n in ("g", "gh", "pool") + post_names + pre_names:
        assert_close_rel(da[n].grad, ra[n].grad, 1e-4, floor=1e-6 * float(ra[n].grad.abs().max()) * ra[n].grad.numel() ** 0.5, what=f"a.d{n}")
    for n in ("pool",) + post_names:
        assert_close_rel(db[n].grad, rb[n].grad, 1e-4, floor=1e-6 * float(rb[n].grad.abs().max()) * rb[n].grad.numel() ** 0.5, what=f"b.d{n}")
