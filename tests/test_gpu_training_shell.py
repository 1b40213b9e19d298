"""GPU: the callers on either side of the hot path: the training-loop shell (utils.pretrain, f-1), the data
set as a drop-in under the reference's DataLoader recipe, and the on-device random draws (f-3)."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import data_oracle

pytestmark = pytest.mark.gpu

TINY = dict(sequences_length=64, num_annotations=301, local_dim=32, global_dim=64, key_dim=16, num_heads=4, num_blocks=2)


@pytest.fixture(scope="module")
def pb():
    import proteinbert_b200
    return proteinbert_b200


def random_samples(n, A, max_len, seed):
    """dummy_tests.create_random_samples (dummy_tests.py:23-38): python `random`, residues uniform over 22 letters, labels ~ B(0.005)."""
    import pandas as pd
    rnd = random.Random(seed)
    rows = []
    for _ in range(n):
        length = rnd.randint(0, max_len)
        seq = "".join(rnd.choice(data_oracle.ALL_AMINO_ACIDS) for _ in range(length))
        ann = [1 if rnd.random() < 0.005 else 0 for _ in range(A)]
        rows.append((seq, ann))
    return pd.DataFrame(rows, columns=["seq", "annotations"])


def test_device_draws_have_the_reference_distributions_and_feed_the_same_kernel(pb):
    dp = pb.data_processing
    rnd = random.Random(1)
    B, L, A = 64, 512, 8943
    seqs = ["".join(rnd.choice(data_oracle.ALL_AMINO_ACIDS) for _ in range(rnd.randint(0, 2 * L))) for _ in range(B)]
    anns = (np.random.RandomState(0).rand(B, A) < 0.005).astype(np.int64)
    (X, Y, W), d = dp.corrupt_pad_device_draws(seqs, anns, L, seed=7, counter=1)
    torch.cuda.synchronize()
    keep, add = d["keep"].cpu().numpy(), d["add"].cpu().numpy()
    tok_mask, tok_rand = d["tok_mask"].cpu().numpy(), d["tok_rand"].cpu().numpy()
    coin, crop = d["coin"].cpu().numpy(), d["crop_start"].cpu().numpy()
    n = keep.size
    assert abs(keep.mean() - 0.75) < 5 * (0.75 * 0.25 / n) ** 0.5                    # multinomial([.25, .75]) -> 1 w.p. 0.75
    assert abs(add.mean() - 1e-4) < 5 * (1e-4 / n) ** 0.5 + 1e-6                      # multinomial([1-1e-4, 1e-4])
    assert abs(tok_mask.mean() - 0.05) < 5 * (0.05 * 0.95 / tok_mask.size) ** 0.5     # multinomial([.95, .05])
    assert tok_rand.min() == 3 and tok_rand.max() == 25                               # randint(3, 26)
    counts = np.bincount(tok_rand, minlength=26)[3:26]
    assert counts.min() > 0.8 * tok_rand.size / 23 and counts.max() < 1.2 * tok_rand.size / 23
    assert (coin >= 0).all() and (coin < 1).all() and 0.25 < coin.mean() < 0.75
    for b, s in enumerate(seqs):                                                      # randint(0, len + 2 - L)
        span = len(s) + 2 - L
        assert 0 <= crop[b] < max(span, 1)
    assert len({int(c) for c in crop[[len(s) + 2 - L > 50 for s in seqs]]}) > 3
    # the kernel output given these draws is what the numpy oracle makes of them (bit-exact), as in the host-draw mode
    off = d["draw_off"].cpu().numpy()
    nd = [dict(crop_start=int(crop[b]), tok_mask=tok_mask[off[b]:off[b + 1]].astype(np.int64), tok_rand=tok_rand[off[b]:off[b + 1]].astype(np.int64),
               coin=float(coin[b]), keep=keep[b].astype(np.int64), add=add[b].astype(np.int64)) for b in range(B)]
    ref = data_oracle.corrupt_batch(seqs, anns, L, nd)
    for got, r in zip((X["local"], X["global"], Y["local"], Y["global"], W["local"], W["global"]), ref):
        assert np.array_equal(got.cpu().numpy(), r)
    # counter-based: same (seed, counter) -> same draws; another counter or seed -> other draws
    _, d2 = dp.corrupt_pad_device_draws(seqs, anns, L, seed=7, counter=1)
    _, d3 = dp.corrupt_pad_device_draws(seqs, anns, L, seed=7, counter=2)
    _, d4 = dp.corrupt_pad_device_draws(seqs, anns, L, seed=8, counter=1)
    assert torch.equal(d2["keep"], d["keep"]) and torch.equal(d2["tok_rand"], d["tok_rand"])
    assert not torch.equal(d3["keep"], d["keep"]) and not torch.equal(d4["keep"], d["keep"])
    # the reference's H5 record arrays as h5py returns them (uniref_dataset.py:236-245): bytes sequences, bool mask rows
    seqs_b = np.array([s.encode("ascii") for s in seqs], dtype=object)
    (Xb, Yb, Wb), _ = dp.corrupt_pad_device_draws(seqs_b, anns.astype(bool), L, seed=7, counter=1)
    for a, b_ in zip((X["local"], X["global"], Y["local"], Y["global"], W["local"], W["global"]),
                     (Xb["local"], Xb["global"], Yb["local"], Yb["global"], Wb["local"], Wb["global"])):
        assert torch.equal(a, b_)


def test_dataset_works_under_the_reference_dataloader_recipe(pb):
    """dummy_tests.py:102-108: DataLoader(dataset, shuffle=True, num_workers=0, batch_size=...) with the DEFAULT collate."""
    from torch.utils.data import DataLoader
    dp = pb.data_processing
    A, L = TINY["num_annotations"], TINY["sequences_length"]
    df = random_samples(10, A, 100, seed=1)
    ds = dp.UniRefGO_PretrainingDataset(df, seq_max_length=L)
    torch.manual_seed(3)
    slow = list(DataLoader(ds, shuffle=False, num_workers=0, batch_size=4))
    torch.manual_seed(3)
    fast = list(DataLoader(ds.raw(), shuffle=False, num_workers=0, batch_size=4, collate_fn=ds.collate))
    assert len(slow) == len(fast) == 3
    for (Xs, Ys, Ws), (Xf, Yf, Wf) in zip(slow, fast):
        for a, b in ((Xs, Xf), (Ys, Yf), (Ws, Wf)):
            assert a["local"].dtype == b["local"].dtype and torch.equal(a["local"], b["local"])
            assert a["global"].dtype == b["global"].dtype and torch.equal(a["global"], b["global"])
    X, Y, W = slow[0]
    assert X["local"].dtype == torch.int64 and X["global"].dtype == torch.float32 and W["local"].dtype == torch.float64
    model = pb.ProteinBERT(device="cuda", **TINY)
    pl, pg = model(X)
    assert pl.shape == (4, L, 26) and pg.shape == (4, A)
    dsd = dp.UniRefGO_PretrainingDataset(df, seq_max_length=L, draws="device", seed=5)
    Xd, Yd, Wd = next(iter(DataLoader(dsd.raw(), batch_size=4, collate_fn=dsd.collate)))
    assert torch.equal(Yd["local"] != 0, Wd["local"] != 0) and Xd["global"].shape == (4, A)


def test_pretrain_shell_checkpoints_resumes_and_pickles(pb, tmp_path):
    """utils.pretrain's contract (utils.py:220-345): results dictionary, checkpoint dictionary, resume, whole-model pickle."""
    from torch.utils.data import DataLoader
    dp = pb.data_processing
    A, L = TINY["num_annotations"], TINY["sequences_length"]
    df = random_samples(16, A, 100, seed=2)
    ds = dp.UniRefGO_PretrainingDataset(df, seq_max_length=L)
    loader = DataLoader(ds.raw(), shuffle=True, batch_size=4, collate_fn=ds.collate)
    torch.manual_seed(0)
    model = pb.ProteinBERT(device="cuda", **TINY)
    opt = pb.FusedAdam(model.parameters(), lr=2e-3)
    w0 = model.pretraining_global_output[0].weight.detach().clone()
    res = pb.pretrain(model, loader, opt, torch.nn.CrossEntropyLoss(reduction="none"), torch.nn.BCELoss(reduction="none"),
                      max_batch_iterations=10, save_path=str(tmp_path), nb_iterations_checkpoint=4, warmup_duration=3)
    assert len(res["train_loss"]) == 10 and all(np.isfinite(res["train_loss"]))
    assert not torch.equal(model.pretraining_global_output[0].weight, w0)
    cks = sorted(f for f in os.listdir(tmp_path) if "checkpoint" in f)
    assert cks == ["proteinbert_pretraining_checkpoint_4.pt", "proteinbert_pretraining_checkpoint_8.pt"]
    ck = torch.load(tmp_path / cks[1], weights_only=False)
    assert set(ck) == {"current_batch_iteration", "model_state_dict", "optimizer_state_dict", "scheduler_state_dict",
                       "warmup_scheduler_state_dict", "full_scheduler_state_dict", "loss"}
    assert ck["current_batch_iteration"] == 8 and len(ck["model_state_dict"]) == len(model.state_dict())
    # whole-model pickle (utils.py:341) loads and computes the same function
    pk = [f for f in os.listdir(tmp_path) if "pretrained_model" in f]
    assert len(pk) == 1
    loaded = torch.load(tmp_path / pk[0], weights_only=False)
    X, _, _ = next(iter(loader))
    with torch.no_grad():
        a, b = model(X), loaded(X)
    assert torch.allclose(a[0], b[0], atol=1e-6) and torch.allclose(a[1], b[1], atol=1e-6)
    # resume: a fresh model + optimizer continue from iteration 8 to 12
    torch.manual_seed(1)
    model2 = pb.ProteinBERT(device="cuda", **TINY)
    opt2 = pb.FusedAdam(model2.parameters(), lr=2e-3)
    res2 = pb.pretrain(model2, loader, opt2, None, None, max_batch_iterations=12, save_path=str(tmp_path), nb_iterations_checkpoint=100,
                       warmup_duration=3, loaded_checkpoint=ck)
    assert len(res2["train_loss"]) == 4
    assert int(next(iter(opt2.state_dict()["state"].values()))["step"]) == 12
