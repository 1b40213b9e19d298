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


def test_gradient_bucket_layout_follows_production_order():
    """training.FlatGradBucket: one flat buffer, gradients laid out in the order backward produces them -- heads,
    blocks last to first (post half, local track, pre half), annotation-input layer and embedding last -- so that everything
    above a cut of backward is a PREFIX of the buffer.  Host logic only: built on CPU parameters."""
    import proteinbert_b200 as pb
    from proteinbert_b200 import training
    cfg = dict(sequences_length=32, num_annotations=50, local_dim=16, global_dim=32, key_dim=8, num_heads=2, num_blocks=4)
    model = pb.ProteinBERT(device="cpu", **cfg)
    bk = training.FlatGradBucket(model)
    names = [n for n, _ in model.named_parameters()]
    assert set(bk.slices) == set(names)
    spans = sorted((off, off + n, name) for name, (off, n) in bk.slices.items())
    for (a0, a1, na), (b0, b1, nb) in zip(spans, spans[1:]):
        assert a1 <= b0, (na, nb)                                   # no overlap
    assert all(off % 4 == 0 for off, _, _ in spans)                 # 16-byte aligned slices (vector stores, multimem)
    assert bk.numel % bk.ALIGN == 0 and spans[-1][1] <= bk.numel
    # group order: heads first, then blocks 3, 2, 1, 0, then the tail
    first = {}
    for off, _, name in spans:
        key = "heads" if name.startswith("pretraining_") else (int(name.split(".")[1]) if name.startswith("proteinBERT_blocks.") else "tail")
        first.setdefault(key, off)
    assert first["heads"] < first[3] < first[2] < first[1] < first[0] < first["tail"]
    # inside a block: post half | local track | pre half
    b2 = [(off, name) for off, _, name in spans if name.startswith("proteinBERT_blocks.2.")]
    post = [o for o, n in b2 if training._is_block_post_param(n)]
    pre = [o for o, n in b2 if training._is_block_pre_param(n)]
    local = [o for o, n in b2 if not training._is_block_post_param(n) and not training._is_block_pre_param(n)]
    assert max(post) < min(local) and max(local) < min(pre)
    # accumulated gradients (biases, LayerNorm affines, W_parameter) are exactly what the clearing launch covers
    covered = lambda off, n: any(a <= off and off + n <= b for a, b in bk.zero_ranges)
    for name, (off, n) in bk.slices.items():
        assert covered(off, n) == training._needs_zero(name) or (not training._needs_zero(name) and not covered(off, n)), name
    # cuts: everything produced above the input of block c is a prefix; cutting lower covers more; with cluster units the post
    # half of the block below belongs to the part above the cut
    o3, o2, o1 = (bk.split_offset(c, units=False) for c in (3, 2, 1))
    assert 0 < o3 < o2 < o1 <= bk.numel
    for c in (3, 2, 1):
        off_u, off_p = bk.split_offset(c, units=True), bk.split_offset(c, units=False)
        assert off_u > off_p
        below_post = [bk.slices[n] for n in names if n.startswith(f"proteinBERT_blocks.{c - 1}.") and training._is_block_post_param(n)]
        assert all(o + k <= off_u for o, k in below_post) and all(o >= off_p for o, k in below_post)
        above = [bk.slices[n] for n in names if n.startswith("pretraining_") or
                 (n.startswith("proteinBERT_blocks.") and int(n.split(".")[1]) >= c)]
        assert all(o + k <= off_p for o, k in above)
    # views alias the flat buffer
    v = bk.view(names[0])
    assert v.data_ptr() == bk.flat.data_ptr() + 4 * bk.slices[names[0]][0] and tuple(v.shape) == tuple(dict(model.named_parameters())[names[0]].shape)


def test_nvls_launch_shape_word():
    """FlatGradBucket.nvls_shape packs (CTAs, threads, reductions in flight) as pb_multimem_allreduce reads them
    (include/pb_b200.h: bits 0-15, 16-27, 28-30 with selector 0 -> 8, 1 -> 1, 2 -> 2, 3 -> 4)."""
    from proteinbert_b200 import training
    w = training.FlatGradBucket.nvls_shape(16, 128, 4)
    assert w & 0xffff == 16 and (w >> 16) & 0xfff == 128 and (w >> 28) & 7 == 3
    assert training.FlatGradBucket.nvls_shape(16) == 16 | (512 << 16)          # defaults: 512 threads, 8 in flight (selector 0)
    assert (training.FlatGradBucket.nvls_shape(148, 32, 2) >> 28) & 7 == 2
    with pytest.raises(KeyError):
        training.FlatGradBucket.nvls_shape(8, 256, 3)
