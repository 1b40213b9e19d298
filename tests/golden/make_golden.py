#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by RUNNING THE REAL REFERENCE.

Run in the build container only (it needs /root/reference, which does not exist
on the GPU box):

    python tests/golden/make_golden.py [--reference /root/reference]

* model_tiny.npz  -- reference ``ProteinBERT/modules.py`` (imports only torch) on a
  tiny configuration: inputs, every registered weight + the unregistered head
  tensors, outputs, the utils.py:293-294 loss, and all gradients, in fp32 and in
  fp64 (the fp64 run calls ``.double()`` and converts the head tensors by hand,
  because they are not registered -- modules.py:73-81).
* model_tiny_seq.npz -- same weights, reference patched in one place
  (``Softmax(dim=1)`` -> ``dim=2`` in GlobalAttentionHead) = the sequence-axis
  softmax variant; outputs only.
* data_path.npz   -- reference ``ProteinBERT/data_processing.py`` run UNMODIFIED
  against a stub of the torchtext primitives it imports (torchtext / h5py are not
  installed here), with every RNG call recorded, for a set of samples covering
  empty, short, exactly-fitting and cropped sequences and out-of-alphabet letters.

Nothing from the reference is copied into this repository; only its outputs.
"""
import argparse
import random
import sys
import types
from collections import OrderedDict

import numpy as np
import torch

TINY = dict(sequences_length=96, num_annotations=301, local_dim=32, global_dim=64,
            key_dim=16, num_heads=4, num_blocks=2)
TINY_B = 4


def head_key(i, h, n):
    return f"proteinBERT_blocks.{i}.global_attention_layer.global_attention_heads.{h}.{n}_parameter"


def all_tensors(model):
    out = OrderedDict((k, v) for k, v in model.named_parameters())
    for i, blk in enumerate(model.proteinBERT_blocks):
        for h, head in enumerate(blk.global_attention_layer.global_attention_heads):
            for n in ("Wq", "Wk", "Wv"):
                out[head_key(i, h, n)] = getattr(head, f"{n}_parameter")
    return out


def make_batch(B, L, A, seed):
    g = torch.Generator().manual_seed(seed)
    lengths = torch.tensor([L, L // 2, 5, L - 1][:B] + [L] * max(0, B - 4))
    pos = torch.arange(L).unsqueeze(0)
    valid = pos < lengths.unsqueeze(1)
    y_local = torch.randint(4, 26, (B, L), generator=g)
    y_local[:, 0] = 1
    y_local[torch.arange(B), lengths - 1] = 2
    y_local = torch.where(valid, y_local, torch.zeros_like(y_local))
    corrupt = (torch.rand(B, L, generator=g) < 0.05) & (y_local > 2)
    x_local = torch.where(corrupt, torch.randint(3, 26, (B, L), generator=g), y_local)
    y_global = (torch.rand(B, A, generator=g) < 0.02).float()
    y_global[2] = 0.0                                  # a sample without annotations -> weight 0
    keep = (torch.rand(B, A, generator=g) < 0.75).float()
    add = (torch.rand(B, A, generator=g) < 0.01).float()
    x_global = (y_global + add) * keep
    x_global[1] = 0.0                                  # hidden annotations branch
    w_local = (y_local != 0).double()
    w_global = y_global.bool().any(dim=1, keepdim=True).double().expand(B, A).contiguous()
    return x_local, x_global, y_local, y_global, w_local, w_global


def run_reference(model, batch):
    import warnings
    xl, xg, yl, yg, wl, wg = batch
    tensors = all_tensors(model)
    for t in tensors.values():
        t.grad = None
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        p_local, p_global = model({"local": xl, "global": xg.to(next(model.parameters()).dtype)})
    # utils.py:293-294 with dummy_tests.py:132-133
    ce = torch.nn.CrossEntropyLoss(reduction="none")
    bce = torch.nn.BCELoss(reduction="none")
    loss = torch.mean(ce(p_local.permute(0, 2, 1), yl) * wl) + \
        torch.mean(bce(p_global, yg.to(p_global.dtype)) * wg)
    loss.backward()
    grads = OrderedDict((k, (t.grad if t.grad is not None else torch.zeros_like(t)).detach().clone())
                        for k, t in tensors.items())
    return p_local.detach(), p_global.detach(), loss.detach(), grads


def gen_model(ref_root, out_dir):
    sys.path.insert(0, f"{ref_root}/ProteinBERT")
    import modules as ref_modules
    torch.manual_seed(1234)
    model = ref_modules.ProteinBERT(device="cpu", **TINY)
    # make the LayerNorm affines non-trivial (the defaults 1/0 would hide indexing bugs)
    g = torch.Generator().manual_seed(99)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if "norm" in name and name.endswith("weight"):
                p.add_(0.2 * torch.randn(p.shape, generator=g))
            if "norm" in name and name.endswith("bias"):
                p.add_(0.1 * torch.randn(p.shape, generator=g))
    batch = make_batch(TINY_B, TINY["sequences_length"], TINY["num_annotations"], seed=5)

    p_local, p_global, loss, grads = run_reference(model, batch)
    store = {}
    for k, t in all_tensors(model).items():
        store["w/" + k] = t.detach().numpy().copy()
    for k, t in grads.items():
        store["g32/" + k] = t.numpy()
    for n, t in zip(("x_local", "x_global", "y_local", "y_global", "w_local", "w_global"), batch):
        store["in/" + n] = t.numpy()
    store["out32/p_local"] = p_local.numpy()
    store["out32/p_global"] = p_global.numpy()
    store["out32/loss"] = loss.numpy()

    # fp64 run of the same reference code
    model.double()
    for blk in model.proteinBERT_blocks:
        for head in blk.global_attention_layer.global_attention_heads:
            for n in ("Wq", "Wk", "Wv"):
                p = getattr(head, f"{n}_parameter")
                p.data = p.data.double()
                p.grad = None
    p_local, p_global, loss, grads = run_reference(model, batch)
    for k, t in grads.items():
        store["g64/" + k] = t.numpy()
    store["out64/p_local"] = p_local.numpy()
    store["out64/p_global"] = p_global.numpy()
    store["out64/loss"] = loss.numpy()
    for k, v in TINY.items():
        store["cfg/" + k] = np.int64(v)
    np.savez_compressed(f"{out_dir}/model_tiny.npz", **store)

    # sequence-axis softmax variant: one-place patch of the reference head
    for blk in model.proteinBERT_blocks:
        for head in blk.global_attention_layer.global_attention_heads:
            head.softmax = torch.nn.Softmax(dim=2)
    p_local, p_global, loss, grads = run_reference(model, batch)
    np.savez_compressed(f"{out_dir}/model_tiny_seq.npz",
                        p_local=p_local.numpy(), p_global=p_global.numpy(), loss=loss.numpy(),
                        **{"g64/" + k: t.numpy() for k, t in grads.items()})
    print("model goldens written; fp64 loss", float(store["out64/loss"]))


def install_torchtext_stub():
    """Minimal stand-ins for the torchtext / h5py names data_processing.py imports
    (behaviour per the torchtext 0.18 documentation)."""
    tt = types.ModuleType("torchtext")
    vocab_mod = types.ModuleType("torchtext.vocab")
    tr = types.ModuleType("torchtext.transforms")

    class Vocab:
        def __init__(self, itos):
            self.itos = list(itos)
            self.stoi = {s: i for i, s in enumerate(self.itos)}
            self.default = None

        def set_default_index(self, i):
            self.default = i

        def __getitem__(self, tok):
            return self.stoi.get(tok, self.default)

        def __len__(self):
            return len(self.itos)

        def get_itos(self):
            return self.itos

        def lookup_indices(self, toks):
            return [self[t] for t in toks]

    def vocab(ordered_dict, min_freq=1, specials=None, special_first=True):
        toks = [t for t, f in ordered_dict.items() if f >= min_freq]
        specials = list(specials or [])
        return Vocab(specials + toks if special_first else toks + specials)

    class Sequential(torch.nn.Sequential):
        def forward(self, x):
            for m in self:
                x = m(x)
            return x

    class VocabTransform(torch.nn.Module):
        def __init__(self, vocab):
            super().__init__()
            self.vocab = vocab

        def forward(self, x):
            return self.vocab.lookup_indices(x)

    class AddToken(torch.nn.Module):
        def __init__(self, token, begin=True):
            super().__init__()
            self.token, self.begin = token, begin

        def forward(self, x):
            return [self.token] + list(x) if self.begin else list(x) + [self.token]

    class ToTensor(torch.nn.Module):
        def __init__(self, padding_value=None, dtype=torch.long):
            super().__init__()
            self.dtype = dtype

        def forward(self, x):
            return torch.tensor(x, dtype=self.dtype)

    class PadTransform(torch.nn.Module):
        def __init__(self, max_length, pad_value):
            super().__init__()
            self.max_length, self.pad_value = max_length, float(pad_value)

        def forward(self, x):
            n = x.size(-1)
            if n < self.max_length:
                x = torch.nn.functional.pad(x, (0, self.max_length - n), value=self.pad_value)
            return x

    vocab_mod.Vocab, vocab_mod.vocab = Vocab, vocab
    for c in (Sequential, VocabTransform, AddToken, ToTensor, PadTransform):
        setattr(tr, c.__name__, c)
    tt.vocab, tt.transforms = vocab_mod, tr
    sys.modules.update({"torchtext": tt, "torchtext.vocab": vocab_mod, "torchtext.transforms": tr,
                        "h5py": types.ModuleType("h5py")})


def gen_data(ref_root, out_dir):
    install_torchtext_stub()
    sys.path.insert(0, f"{ref_root}/ProteinBERT")
    import pandas as pd
    import data_processing as ref_dp

    L, A = 64, 157
    letters = "ACDEFGHIKLMNPQRSTUVWXY"
    rnd = random.Random(7777)

    def rseq(n, alphabet=letters):
        return "".join(alphabet[rnd.randint(0, len(alphabet) - 1)] for _ in range(n))

    seqs = ["", "A", rseq(10), rseq(L - 3), rseq(L - 2), rseq(L - 1), rseq(L), rseq(L + 1),
            rseq(3 * L), rseq(40, letters + "BJOZ*x"), rseq(200), rseq(61), rseq(62), rseq(63)]
    anns = []
    for i in range(len(seqs)):
        if i in (2, 5):
            anns.append([0] * A)                      # no annotation at all -> weight 0
        else:
            anns.append([1 if rnd.random() < 0.05 else 0 for _ in range(A)])
    df = pd.DataFrame(list(zip(seqs, anns)))
    ds = ref_dp.UniRefGO_PretrainingDataset(df, seq_max_length=L)

    # record every RNG call the reference makes, in order
    log = []
    orig = dict(randint=torch.randint, multinomial=torch.multinomial, rand=torch.rand)

    def wrap(name):
        def f(*a, **k):
            r = orig[name](*a, **k)
            log.append((name, r.clone()))
            return r
        return f

    torch.manual_seed(0)
    store = {"L": np.int64(L), "A": np.int64(A)}
    # two passes with different RNG streams so both annotation branches show up per sample
    rec = 0
    for rep in range(3):
        for i in range(len(seqs)):
            log.clear()
            for n in orig:
                setattr(torch, n, wrap(n))
            try:
                X, Y, W = ds[i]
            finally:
                for n, f in orig.items():
                    setattr(torch, n, f)
            names = [n for n, _ in log]
            k = 0
            crop = -1
            n_ids = len(seqs[i]) + 2
            if n_ids > L:
                assert names[k] == "randint"
                crop = int(log[k][1][0]); k += 1
            assert names[k] == "multinomial"; tok_mask = log[k][1].numpy(); k += 1
            assert names[k] == "randint"; tok_rand = log[k][1].numpy(); k += 1
            assert names[k] == "rand"; coin = float(log[k][1][0]); k += 1
            if coin > 0.5:
                keep = log[k][1].numpy(); add = log[k + 1][1].numpy(); k += 2
            else:
                keep = np.zeros(A, np.int64); add = np.zeros(A, np.int64)
            assert k == len(log), (names, k)
            pre = f"s{rec}/"
            store[pre + "seq"] = np.frombuffer(seqs[i].encode("latin-1"), dtype=np.uint8).copy()
            store[pre + "ann"] = np.asarray(anns[i], dtype=np.int64)
            store[pre + "crop"] = np.int64(crop)
            store[pre + "tok_mask"] = tok_mask.astype(np.int64)
            store[pre + "tok_rand"] = tok_rand.astype(np.int64)
            store[pre + "coin"] = np.float32(coin)
            store[pre + "keep"] = keep.astype(np.int64)
            store[pre + "add"] = add.astype(np.int64)
            store[pre + "x_local"] = X["local"].numpy()
            store[pre + "x_global"] = X["global"].numpy()
            store[pre + "y_local"] = Y["local"].numpy()
            store[pre + "y_global"] = Y["global"].numpy()
            store[pre + "w_local"] = np.asarray(W["local"])
            store[pre + "w_global"] = np.asarray(W["global"])
            assert X["local"].dtype == torch.int64 and X["global"].dtype == torch.float32
            assert store[pre + "w_local"].dtype == np.float64 and store[pre + "w_global"].dtype == np.float64
            rec += 1
    store["n"] = np.int64(rec)
    vocab = ref_dp.create_amino_acid_vocab()
    store["itos"] = np.array(vocab.get_itos())
    np.savez_compressed(f"{out_dir}/data_path.npz", **store)
    print("data goldens written:", rec, "records")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--out", default=__file__.rsplit("/", 1)[0])
    a = ap.parse_args()
    gen_model(a.reference, a.out)
    gen_data(a.reference, a.out)
