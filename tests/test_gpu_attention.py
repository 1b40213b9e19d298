"""GPU: the attention layer as the reference EXECUTES it (csrc/ops_attention.cu; modules.py:21-92) and the sequence-axis mode
 against the float64 oracles -- attention_as_executed, attention_closed_form (the two must agree) and
attention_sequence -- values and every gradient, at a small odd shape and at the default model's shape; then whole-model
steps in softmax_axis="sequence" against the golden file a dim=2-patched copy of the real reference produced
(tests/golden/model_tiny_seq.npz).  fp32 engine: 1e-4 relative, absolute floor for mathematically-zero gradients."""
import numpy as np
import pytest
import torch

from helpers import GOLDEN, assert_close_rel, golden_batch, golden_params, load_model_golden
from oracle import model_oracle
from test_gpu_parity import build_model, run_step

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pb():
    import proteinbert_b200
    return proteinbert_b200


def case(B, L, C, G, K, H, seed):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s, scale=1.0: torch.randn(*s, generator=g) * scale
    Dv = G // H
    return dict(h=r(B, L, C), g=r(B, G), Wq=r(H, G, K, scale=G ** -0.5), Wk=r(H, C, K, scale=C ** -0.5), Wv=r(H, C, Dv, scale=C ** -0.5),
                Wp=r(K), datt=r(B, G, scale=0.1))


@pytest.mark.parametrize("B,L,C,G,K,H", [(3, 40, 32, 64, 16, 4), (2, 512, 128, 512, 64, 4), (2, 70, 48, 96, 8, 3)])
@pytest.mark.parametrize("axis", ["reference_as_executed", "sequence"])
def test_fused_attention_against_float64_oracles(pb, B, L, C, G, K, H, axis):
    from proteinbert_b200 import ops
    t = case(B, L, C, G, K, H, seed=L + K)
    dev = {k: v.cuda().requires_grad_(k != "datt") for k, v in t.items()}
    ref = {k: v.double().requires_grad_(k != "datt") for k, v in t.items()}
    att = ops.AttentionFn.apply(dev["h"], dev["g"], dev["Wq"], dev["Wk"], dev["Wv"], dev["Wp"], 0 if axis != "sequence" else 1)
    heads = [(ref["Wq"][i], ref["Wk"][i], ref["Wv"][i]) for i in range(H)]
    fn = model_oracle.attention_sequence if axis == "sequence" else model_oracle.attention_as_executed
    r_att = fn(ref["h"], ref["g"], heads, ref["Wp"])
    assert_close_rel(att, r_att, 1e-4, what="att")
    if axis != "sequence":      # the closed form the tensor-core path evaluates is the same function
        assert_close_rel(att, model_oracle.attention_closed_form(ref["h"].detach(), ref["g"].detach(), [tuple(x.detach() for x in hd) for hd in heads],
                                                                 ref["Wp"].detach()), 1e-4, what="closed form")
    att.backward(dev["datt"])
    r_att.backward(ref["datt"])
    torch.cuda.synchronize()
    gmax = max(float(ref[n].grad.abs().max()) for n in ("h", "Wv"))
    for n in ("h", "g", "Wq", "Wk", "Wv", "Wp"):
        floor = 1e-5 * gmax * ref[n].grad.numel() ** 0.5          # under the reference axis dg, dWq, dWk are mathematically zero
        assert_close_rel(dev[n].grad, ref[n].grad, 1e-4 if n != "Wp" else 2e-3, floor=floor, what=f"d{n} ({axis})")
    if axis != "sequence":
        assert float(dev["Wq"].grad.abs().max()) == 0.0 and float(dev["Wk"].grad.abs().max()) == 0.0


@pytest.mark.parametrize("mode", ["reference_as_executed", "sequence"])
def test_whole_model_in_both_attention_modes(pb, mode):
    """softmax_axis="sequence" against the golden of the dim=2-patched reference; "reference_as_executed" against the golden of
    the unmodified reference (it must give what the closed-form default gives)."""
    gold = load_model_golden()
    cfg = {k: int(v) for k, v in gold["cfg"].items()}
    model = pb.ProteinBERT(device="cuda", precision="fp32", softmax_axis=mode, **cfg)
    params = golden_params(gold)
    with torch.no_grad():
        for k, v in model.state_dict().items():
            v.copy_(params[k])
        for k, t in model.head_tensors().items():
            t.copy_(params[k])
    p_local, p_global, loss, grads = run_step(pb, model, golden_batch(gold))
    if mode == "sequence":
        z = np.load(GOLDEN + "/model_tiny_seq.npz")
        ref = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("g64/")}
        rp_local, rp_global, rloss = z["p_local"], z["p_global"], float(z["loss"])
    else:
        ref = {k: torch.from_numpy(v) for k, v in gold["g64"].items()}
        rp_local, rp_global, rloss = gold["out64"]["p_local"], gold["out64"]["p_global"], float(gold["out64"]["loss"])
    assert_close_rel(p_local, rp_local, 1e-4, what="p_local")
    assert_close_rel(p_global, rp_global, 1e-4, what="p_global")
    assert abs(float(loss) - rloss) <= 1e-4 * abs(rloss)
    gmax = max(float(v.abs().max()) for v in ref.values())
    named = dict(model.named_parameters())
    heads = model.head_tensors()
    for k, r in ref.items():
        got = named[k].grad if k in named else heads[k].grad
        if got is None:
            got = torch.zeros_like(r)
        zero_like = ("pretraining_local_output.0.bias" in k) or (mode != "sequence" and ("Wq_" in k or "Wk_" in k))
        floor = 1e-6 * gmax * np.sqrt(r.numel()) if (zero_like or k.endswith("W_parameter")) else 0.0
        assert_close_rel(got, r, 2e-3 if k.endswith("W_parameter") else 1e-4, floor=floor, what=f"{mode}/{k}")


def test_register_heads_puts_the_head_tensors_into_the_state_dict(pb):
    cfg = dict(sequences_length=64, num_annotations=30, local_dim=32, global_dim=64, key_dim=16, num_heads=4, num_blocks=2)
    a = pb.ProteinBERT(device="cuda", **cfg)
    b = pb.ProteinBERT(device="cuda", register_heads=True, softmax_axis="sequence", **cfg)
    assert len(a.state_dict()) == 49 and len(b.state_dict()) == 49 + 2 * 4 * 3
    X = {"local": torch.randint(0, 26, (3, 64), device="cuda"), "global": torch.zeros(3, 30, device="cuda")}
    pl, pg = b(X)
    (pl.sum() + pg.sum()).backward()
    assert all(p.grad is not None for p in b.parameters())
