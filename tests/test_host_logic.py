"""CPU: host-side mirror of the reference interface (no kernels are launched)."""
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN, load_model_golden
from oracle import data_oracle, model_oracle


@pytest.fixture(scope="module")
def pb():
    import proteinbert_b200
    return proteinbert_b200


def test_state_dict_keys_and_shapes_match_the_reference(pb):
    gold = load_model_golden()
    cfg = {k: int(v) for k, v in gold["cfg"].items()}
    model = pb.ProteinBERT(device="cpu", **cfg)
    sd = model.state_dict()
    ref_reg = {k: v for k, v in gold["w"].items() if "global_attention_heads" not in k}
    assert list(sd.keys()) == list(ref_reg.keys())          # same keys, same order as the reference
    for k, v in ref_reg.items():
        assert tuple(sd[k].shape) == tuple(v.shape), k
    heads = model.head_tensors()
    assert set(heads) == {k for k in gold["w"] if "global_attention_heads" in k}
    # heads are NOT parameters of the model (plain python list, modules.py:73-81)
    assert len(list(model.parameters())) == len(ref_reg)
    assert isinstance(model.proteinBERT_blocks[0].global_attention_layer.global_attention_heads, list)


def test_default_config_parameter_count(pb):
    model = pb.ProteinBERT(512, 8943, 128, 512, 64, 4, 6, device="cpu")
    assert len(model.state_dict()) == 133
    assert sum(p.numel() for p in model.parameters()) == 16175241


def test_same_seed_same_initial_weights_as_reference_draw_order(pb):
    """Sub-modules are created in the reference's order, so the golden weights (reference, seed 1234,
    LayerNorm affines perturbed afterwards) must be reproduced for every non-LayerNorm tensor."""
    gold = load_model_golden()
    cfg = {k: int(v) for k, v in gold["cfg"].items()}
    torch.manual_seed(1234)
    model = pb.ProteinBERT(device="cpu", **cfg)
    mine = dict(model.state_dict())
    mine.update(model.head_tensors())
    for k, v in gold["w"].items():
        if "norm" in k:
            continue
        assert np.array_equal(mine[k].detach().numpy(), v), k


def test_cpu_tensors_are_refused_not_silently_computed(pb):
    model = pb.ProteinBERT(32, 11, 16, 32, 8, 2, 1, device="cpu")
    x = {"local": torch.zeros(2, 32, dtype=torch.long), "global": torch.zeros(2, 11)}
    with pytest.raises(pb.PBError):
        model(x)


def test_block_rejects_wrong_sequence_length_and_precision(pb):
    with pytest.raises(ValueError):
        pb.ProteinBERT(32, 11, 16, 32, 8, 2, 1, device="cpu", precision="fp16")
    with pytest.raises(AssertionError):
        pb.ProteinBERTBlock(32, 16, 30, 4, 8, device="cpu")


def test_draws_follow_the_reference_rng_order(pb):
    """Same seed, same sample order -> the same draws the real reference made (recorded in the golden)."""
    z = np.load(os.path.join(GOLDEN, "data_path.npz"))
    L, A = int(z["L"]), int(z["A"])
    torch.manual_seed(0)
    for r in range(int(z["n"])):
        pre = f"s{r}/"
        d = pb.data_processing.draw_sample(len(z[pre + "seq"]), A, L)
        assert d["crop_start"] == int(z[pre + "crop"]), r
        assert np.array_equal(d["tok_mask"].numpy(), z[pre + "tok_mask"]), r
        assert np.array_equal(d["tok_rand"].numpy(), z[pre + "tok_rand"]), r
        assert np.float32(d["coin"]) == z[pre + "coin"], r
        if d["keep"] is not None:
            assert np.array_equal(d["keep"].numpy(), z[pre + "keep"]) and np.array_equal(d["add"].numpy(), z[pre + "add"])


def test_pack_batch_layout(pb):
    seqs = ["ACD", "", "W" * 10]
    anns = [[0, 1, 0], [0, 0, 0], [1, 1, 1]]
    torch.manual_seed(3)
    draws = [pb.data_processing.draw_sample(len(s), 3, 6) for s in seqs]
    p = pb.data_processing.pack_batch(seqs, anns, draws, 6, pin=False)
    assert p["seq_off"].tolist() == [0, 3, 3, 13]
    assert p["draw_off"].tolist() == [0, 5, 7, 13]
    assert p["seq_bytes"].numpy().tobytes() == b"ACDWWWWWWWWWW"
    assert p["ann"].dtype == torch.uint8 and p["keep"].shape == (3, 3)
    assert p["crop_start"][2] >= 0 and draws[2]["crop_start"] >= 0 and draws[0]["crop_start"] == -1


def test_vocab_matches_reference(pb):
    v = pb.data_processing.create_amino_acid_vocab()
    z = np.load(os.path.join(GOLDEN, "data_path.npz"))
    assert v.get_itos() == list(z["itos"])
    assert len(v) == 26 and v["<pad>"] == 0 and v["out of vocab"] == v["<unk>"] == 3
    lut = data_oracle.char_lut()
    assert all(lut[ord(c)] == v[c] for c in data_oracle.ALL_AMINO_ACIDS)


def test_reference_state_dict_loads_strictly_and_the_module_pickles(tmp_path):
    """utils.py:270 (load_state_dict of a reference checkpoint) and :341 (torch.save(model)): the golden file carries the REAL
    reference's state_dict (tests/golden/make_golden.py ran the reference); it must load with strict=True, and the module
    must survive a pickle round trip with the same tensors (no GPU needed: the modules are parameter containers)."""
    import torch
    import proteinbert_b200 as pb
    from helpers import load_model_golden
    gold = load_model_golden()
    cfg = {k: int(v) for k, v in gold["cfg"].items()}
    model = pb.ProteinBERT(device="cpu", **cfg)
    sd = {k: torch.from_numpy(v) for k, v in gold["w"].items() if k in model.state_dict()}
    assert len(sd) == len(model.state_dict())
    missing, unexpected = model.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    path = tmp_path / "model.pt"
    torch.save(model, path)
    back = torch.load(path, weights_only=False)
    assert list(back.state_dict()) == list(model.state_dict())
    for k, v in model.state_dict().items():
        assert torch.equal(back.state_dict()[k], v), k
    heads_a, heads_b = model.head_tensors(), back.head_tensors()       # the unregistered head tensors travel with the pickle too
    assert all(torch.equal(heads_a[k], heads_b[k]) for k in heads_a)


def test_warmup_then_plateau_follows_the_reference_schedule():
    """utils.py:253-264: lr = base * step / warmup for `warmup` steps (0 at the very first step), then ReduceLROnPlateau."""
    import torch
    import proteinbert_b200 as pb
    p = torch.nn.Parameter(torch.zeros(3))
    opt = torch.optim.SGD([p], lr=1.0)
    sch = pb.WarmupThenPlateau(opt, warmup_duration=4, patience=2)
    lrs = [sch.get_last_lr()[0]]
    for i in range(4):
        sch.step(1.0)
        lrs.append(sch.get_last_lr()[0])
    assert lrs == [0.0, 0.25, 0.5, 0.75, 1.0]
    for _ in range(4):                       # a flat loss: patience 2 -> the plateau scheduler cuts the rate by 10
        sch.step(1.0)
    assert abs(sch.get_last_lr()[0] - 0.1) < 1e-12
    sd = sch.state_dict()
    assert set(sd) == {"scheduler_state_dict", "warmup_scheduler_state_dict", "full_scheduler_state_dict"}
    opt2 = torch.optim.SGD([p], lr=1.0)
    sch2 = pb.WarmupThenPlateau(opt2, warmup_duration=4, patience=2)
    sch2.load_state_dict(sd)
    assert sch2.steps == sch.steps
