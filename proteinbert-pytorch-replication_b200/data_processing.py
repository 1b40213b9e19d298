"""Host side of the input corruption / padding path (reference: ProteinBERT/data_processing.py:30-183,337-348).

The reference builds every sample in Python on the CPU (tokenise -> <sos>/<eos> -> random crop -> 5 % token
randomisation -> right-pad; annotation noising; weights).  Here the host only makes the RANDOM DRAWS -- with
the same torch CPU RNG calls, shapes and order as the reference, so the same seed gives the same draws -- and
packs the raw bytes; all integer work then runs in one CUDA kernel (``pb_corrupt_pad``) that is bit-exact
with the reference given those draws.

    ds     = UniRefGO_PretrainingDataset(df, seq_max_length=512)       # same constructor as the reference
    loader = DataLoader(ds, batch_size=64)                             # drop-in: one kernel launch per sample
    loader = DataLoader(ds.raw(), batch_size=64, collate_fn=ds.collate)   # fast: one kernel launch per batch
    for X, Y, W in loader: ...                                         # dicts of device tensors

There is no torchtext dependency (the reference's is un-vendored, unpinned and end-of-life).
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import numpy as np
import torch

from ._lib import check, lib

ALL_AMINO_ACIDS = "ACDEFGHIKLMNPQRSTUVWXY"          # data_processing.py:338
SPECIAL_TOKENS = ["<pad>", "<sos>", "<eos>", "<unk>"]  # :339


class AminoAcidVocab:
    """What ``create_amino_acid_vocab()`` returns in the reference, minus torchtext: specials first, then the
    22 residue letters; unknown tokens map to ``<unk>`` (data_processing.py:337-348)."""

    def __init__(self):
        self._itos = SPECIAL_TOKENS + list(ALL_AMINO_ACIDS)
        self._stoi = {s: i for i, s in enumerate(self._itos)}
        self._default = self._stoi["<unk>"]

    def __getitem__(self, token: str) -> int:
        return self._stoi.get(token, self._default)

    def __len__(self) -> int:
        return len(self._itos)

    def get_itos(self) -> List[str]:
        return list(self._itos)


def create_amino_acid_vocab() -> AminoAcidVocab:
    return AminoAcidVocab()


def draw_sample(n_residues: int, num_annotations: int, max_length: int, p: float = .05,
                positive_p: float = 0.25, negative_p: float = 0.0001) -> Dict[str, object]:
    """The random draws ``UniRefGO_PretrainingDataset.__getitem__`` makes for one sample, in its order
    (crop randint -> token multinomial -> replacement randint -> coin rand -> keep/add multinomials;
    data_processing.py:78-83,100-105,125-142) on the torch CPU generator."""
    n_ids = n_residues + 2
    crop = -1
    if n_ids > max_length:
        crop = int(torch.randint(0, n_ids - max_length, (1,))[0])
    n = min(n_ids, max_length)
    tok_mask = torch.multinomial(torch.tensor([1 - p, p], dtype=torch.float), num_samples=n, replacement=True)
    tok_rand = torch.randint(3, 4 + len(ALL_AMINO_ACIDS), (n,))
    coin = float(torch.rand(1)[0])
    if coin > 0.5:
        keep = torch.multinomial(torch.tensor([positive_p, 1 - positive_p], dtype=torch.float),
                                 num_samples=num_annotations, replacement=True)
        add = torch.multinomial(torch.tensor([1 - negative_p, negative_p], dtype=torch.float),
                                num_samples=num_annotations, replacement=True)
    else:
        keep = add = None
    return {"crop_start": crop, "tok_mask": tok_mask, "tok_rand": tok_rand, "coin": coin, "keep": keep, "add": add}


def pack_batch(seqs: Sequence[str], anns, draws: Sequence[Dict[str, object]], max_length: int, pin: bool = True):
    """Ragged python samples + their draws -> flat host tensors in the layout ``pb_corrupt_pad`` reads."""
    B = len(seqs)
    A = len(anns[0])
    raw = [s.encode("latin-1", errors="replace") for s in seqs]
    seq_off = np.zeros(B + 1, dtype=np.int64)
    seq_off[1:] = np.cumsum([len(r) for r in raw])
    seq_bytes = np.frombuffer(b"".join(raw) or b"\0", dtype=np.uint8).copy()
    n_draw = [min(len(r) + 2, max_length) for r in raw]
    draw_off = np.zeros(B + 1, dtype=np.int64)
    draw_off[1:] = np.cumsum(n_draw)
    tok_mask = np.concatenate([np.asarray(d["tok_mask"], dtype=np.int64)[:n] for d, n in zip(draws, n_draw)]).astype(np.uint8)
    tok_rand = np.concatenate([np.asarray(d["tok_rand"], dtype=np.int64)[:n] for d, n in zip(draws, n_draw)]).astype(np.uint8)
    crop = np.asarray([max(int(d["crop_start"]), 0) for d in draws], dtype=np.int64)
    coin = np.asarray([d["coin"] for d in draws], dtype=np.float32)
    keep = np.zeros((B, A), dtype=np.uint8)
    add = np.zeros((B, A), dtype=np.uint8)
    for i, d in enumerate(draws):
        if d["keep"] is not None:
            keep[i] = np.asarray(d["keep"], dtype=np.int64)
            add[i] = np.asarray(d["add"], dtype=np.int64)
    ann = np.asarray(anns, dtype=np.int64)
    if ann.min(initial=0) < 0 or ann.max(initial=0) > 255:
        raise ValueError("annotation values must fit in a byte (the reference uses 0/1 masks)")
    packed = {"seq_bytes": seq_bytes, "seq_off": seq_off, "crop_start": crop, "tok_mask": tok_mask, "tok_rand": tok_rand,
              "draw_off": draw_off, "coin": coin, "keep": keep, "add": add, "ann": ann.astype(np.uint8)}
    out = {}
    for k, v in packed.items():
        t = torch.from_numpy(np.ascontiguousarray(v))
        out[k] = t.pin_memory() if (pin and torch.cuda.is_available()) else t
    out["B"], out["L"], out["A"] = B, max_length, A
    return out


def corrupt_pad(packed, device="cuda", out=None):
    """Run the corruption/padding kernel.  Returns (X, Y, W) dicts of device tensors with the reference's
    shapes and dtypes: local int64 [B,L] / global float32 [B,A]; weights float64.  ``out``: the (X, Y, W) triple of an
    earlier call of the same shape, to be overwritten (a training loop that has consumed the previous batch saves six
    allocations per batch)."""
    B, L, A = packed["B"], packed["L"], packed["A"]
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("corrupt_pad runs on a CUDA device only (no CPU fallback)")
    d = {k: v.to(dev, non_blocking=True) for k, v in packed.items() if isinstance(v, torch.Tensor)}
    if out is not None:
        (X, Y, W) = out
        x_local, y_local, w_local, x_global, y_global, w_global = X["local"], Y["local"], W["local"], X["global"], Y["global"], W["global"]
        want = (((B, L), torch.int64), ((B, L), torch.int64), ((B, L), torch.float64), ((B, A), torch.float32), ((B, A), torch.float32),
                ((B, A), torch.float64))
        for t, (shp, dty) in zip((x_local, y_local, w_local, x_global, y_global, w_global), want):
            same_dev = t.device.type == "cuda" and (dev.index is None or t.device.index == dev.index)
            if tuple(t.shape) != shp or t.dtype != dty or not same_dev or not t.is_contiguous():
                raise ValueError("corrupt_pad(out=...): tensors of another shape, dtype or device")
    else:
        x_local = torch.empty((B, L), dtype=torch.int64, device=dev)
        y_local = torch.empty((B, L), dtype=torch.int64, device=dev)
        w_local = torch.empty((B, L), dtype=torch.float64, device=dev)
        x_global = torch.empty((B, A), dtype=torch.float32, device=dev)
        y_global = torch.empty((B, A), dtype=torch.float32, device=dev)
        w_global = torch.empty((B, A), dtype=torch.float64, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    check(lib().pb_corrupt_pad(d["seq_bytes"].data_ptr(), d["seq_off"].data_ptr(), d["crop_start"].data_ptr(),
                               d["tok_mask"].data_ptr(), d["tok_rand"].data_ptr(), d["draw_off"].data_ptr(),
                               d["coin"].data_ptr(), d["keep"].data_ptr(), d["add"].data_ptr(), d["ann"].data_ptr(),
                               x_local.data_ptr(), y_local.data_ptr(), w_local.data_ptr(), x_global.data_ptr(),
                               y_global.data_ptr(), w_global.data_ptr(), B, L, A, st), "pb_corrupt_pad")
    return ({"local": x_local, "global": x_global}, {"local": y_local, "global": y_global},
            {"local": w_local, "global": w_global})


def corrupt_pad_device_draws(seqs, anns, max_length: int, seed: int, counter: int, device="cuda",
                             p: float = .05, positive_p: float = 0.25, negative_p: float = 0.0001):
    """corrupt_pad with the random draws made on the GPU (``pb_draw_batch``, Philox keyed by ``seed``, ``counter`` = batch
    number) instead of the torch CPU generator: the host only packs the residue bytes and the annotation rows.
    ``seqs`` may be str or bytes (h5py hands the ``seqs`` dataset of the reference's H5 record format back as bytes,
    uniref_dataset.py:236-245) and ``anns`` any integer or bool [B, A] array (its ``annotation_masks`` rows)."""
    B, A = len(seqs), len(anns[0])
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("device draws run on a CUDA device only")
    raw = [bytes(s) if isinstance(s, (bytes, bytearray, np.bytes_)) else str(s).encode("latin-1", errors="replace") for s in seqs]
    n_draw = [min(len(r) + 2, max_length) for r in raw]
    seq_off = np.zeros(B + 1, dtype=np.int64); seq_off[1:] = np.cumsum([len(r) for r in raw])
    draw_off = np.zeros(B + 1, dtype=np.int64); draw_off[1:] = np.cumsum(n_draw)
    seq_bytes = np.frombuffer(b"".join(raw) or b"\0", dtype=np.uint8).copy()
    ann = np.asarray(anns, dtype=np.int64)
    if ann.min(initial=0) < 0 or ann.max(initial=0) > 255:
        raise ValueError("annotation values must fit in a byte (the reference uses 0/1 masks)")
    h = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().to(dev, non_blocking=True)
    d = {"seq_bytes": h(seq_bytes), "seq_off": h(seq_off), "draw_off": h(draw_off), "ann": h(ann.astype(np.uint8)),
         "crop_start": torch.empty(B, dtype=torch.int64, device=dev),
         "tok_mask": torch.empty(max(int(draw_off[-1]), 1), dtype=torch.uint8, device=dev),
         "tok_rand": torch.empty(max(int(draw_off[-1]), 1), dtype=torch.uint8, device=dev),
         "coin": torch.empty(B, dtype=torch.float32, device=dev),
         "keep": torch.empty((B, A), dtype=torch.uint8, device=dev), "add": torch.empty((B, A), dtype=torch.uint8, device=dev),
         "B": B, "L": max_length, "A": A}
    st = torch.cuda.current_stream(dev).cuda_stream
    check(lib().pb_draw_batch(int(seed) & (2 ** 64 - 1), int(counter) & (2 ** 64 - 1), d["seq_off"].data_ptr(), d["draw_off"].data_ptr(),
                              d["crop_start"].data_ptr(), d["tok_mask"].data_ptr(), d["tok_rand"].data_ptr(), d["coin"].data_ptr(),
                              d["keep"].data_ptr(), d["add"].data_ptr(), B, max_length, A, float(p), float(1.0 - positive_p),
                              float(negative_p), st), "pb_draw_batch")
    return corrupt_pad(d, dev), d


class UniRefGO_PretrainingDataset(torch.utils.data.Dataset):
    """Same constructor, length and per-sample output as the reference dataset (data_processing.py:146-183): ``ds[i]`` is the
    ``(X, Y, W)`` triple of dictionaries -- ``local`` int64 [L] / ``global`` float32 [A], weights float64 -- so the reference's
    ``DataLoader(ds, shuffle=True, num_workers=0, batch_size=32)`` (dummy_tests.py:102-108) works unchanged; the tensors are on
    the GPU already (each sample is one ``pb_corrupt_pad`` launch of batch 1, default_collate stacks them there).
    Iterating with the same seed visits the torch CPU RNG exactly as the reference does, so the same seed gives the same batch.

    Faster: ``DataLoader(ds.raw(), batch_size=32, collate_fn=ds.collate)`` hands the raw samples and their draws to ONE kernel
    launch per batch.  ``draws="device"`` makes ``collate`` take the random draws from the Philox generator on the GPU
    (``pb_draw_batch``) instead of the torch CPU generator: the same distributions, not the same stream -- for training
    runs where the host draws (two 8943-element multinomials per sample) would be the bottleneck."""

    def __init__(self, df, seq_max_length=128, device="cuda", draws="host", seed=0):
        if draws not in ("host", "device"):
            raise ValueError("draws must be 'host' (the reference's torch CPU RNG stream) or 'device' (Philox on the GPU)")
        self.df = df
        self.vocab = create_amino_acid_vocab()
        self.seq_max_length = seq_max_length
        self.device = device
        self.draws = draws
        self.seed = int(seed)
        self._batches = 0

    def raw_item(self, index):
        seq = self.df.iloc[index, 0]
        ann = self.df.iloc[index, 1]
        return seq, ann, (draw_sample(len(seq), len(ann), self.seq_max_length) if self.draws == "host" else None)

    def __getitem__(self, index):
        seq, ann, d = self.raw_item(index)
        if d is None:
            d = draw_sample(len(seq), len(ann), self.seq_max_length)
        X, Y, W = corrupt_pad(pack_batch([seq], [ann], [d], self.seq_max_length), self.device)
        return tuple({k: v[0] for k, v in t.items()} for t in (X, Y, W))

    def __len__(self):
        return len(self.df.index)

    def raw(self):
        """A view of this dataset that yields raw samples (sequence, annotations, draws) for ``collate``."""
        return _RawView(self)

    def collate(self, samples):
        seqs, anns, draws = zip(*samples)
        if self.draws == "device":
            self._batches += 1
            return corrupt_pad_device_draws(seqs, anns, self.seq_max_length, self.seed, self._batches, self.device)[0]
        return corrupt_pad(pack_batch(seqs, anns, draws, self.seq_max_length), self.device)


class _RawView(torch.utils.data.Dataset):
    def __init__(self, ds):
        self.ds = ds

    def __getitem__(self, index):
        return self.ds.raw_item(index)

    def __len__(self):
        return len(self.ds)


def synthetic_batch(B: int, L: int, A: int, seed: int = 0, vocab_size: int = 26):
    """Synthetic (X, Y, W) tensors of the shapes/dtypes the reference dataset yields (data_processing.py:159-180) without the
    string path, for benchmarks and probes: random lengths (padded and full-length rows), 5 %-corrupted X tokens,
    Bernoulli(0.005) annotations, half of the samples with all annotations hidden.  CPU tensors."""
    gen = torch.Generator().manual_seed(seed)
    lengths = torch.randint(2, 2 * L, (B,), generator=gen).clamp(max=L)
    pos = torch.arange(L).unsqueeze(0)
    valid = pos < lengths.unsqueeze(1)
    y_local = torch.randint(4, vocab_size, (B, L), generator=gen)
    y_local[:, 0] = 1
    y_local[torch.arange(B), lengths - 1] = 2
    y_local = torch.where(valid, y_local, torch.zeros_like(y_local))
    corrupt = (torch.rand(B, L, generator=gen) < 0.05) & (y_local > 2)
    x_local = torch.where(corrupt, torch.randint(3, vocab_size, (B, L), generator=gen), y_local)
    y_global = (torch.rand(B, A, generator=gen) < 0.005).float()
    keep = (torch.rand(B, A, generator=gen) < 0.75).float()
    add = (torch.rand(B, A, generator=gen) < 1e-4).float()
    shown = (torch.rand(B, 1, generator=gen) > 0.5).float()
    x_global = (y_global + add) * keep * shown
    w_local = (y_local != 0).double()
    w_global = y_global.bool().any(dim=1, keepdim=True).double().expand(B, A).contiguous()
    return x_local, x_global, y_local, y_global, w_local, w_global
