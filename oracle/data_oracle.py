"""CPU oracle for the input corruption / padding path (integer work, numpy).

TEST INFRASTRUCTURE ONLY -- imported by ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s CPU-baseline leg, never by the product package.

Restates ``UniRefGO_PretrainingDataset.__getitem__`` (data_processing.py:159-180)
and the transforms it calls, as a pure function of the sample and of the random
DRAWS the reference makes (in its order: crop ``randint``, token ``multinomial``,
replacement ``randint``, annotation coin ``rand``, keep ``multinomial``, add
``multinomial``).  Given the same draws the result must be bit-identical; the
goldens in ``tests/golden/data_path.npz`` were produced by running the real
``data_processing.py`` (against a stub of the six torchtext primitives it uses)
with the RNG calls recorded -- see ``tests/golden/make_golden.py``.

The torchtext primitives themselves are a third-party dependency that the
reference neither vendors nor pins (project EOL at 0.18); their documented
behaviour is restated here: ``VocabTransform`` = table lookup with a default
index, ``AddToken`` = prepend / append, ``ToTensor`` = int64 tensor,
``PadTransform`` = right-pad the last dimension up to ``max_length``.
"""
from __future__ import annotations

import numpy as np

# data_processing.py:337-348 -- specials first, then the 22 residue letters.
ALL_AMINO_ACIDS = "ACDEFGHIKLMNPQRSTUVWXY"
PAD, SOS, EOS, UNK = 0, 1, 2, 3
VOCAB_SIZE = 4 + len(ALL_AMINO_ACIDS)          # 26


def char_lut() -> np.ndarray:
    """256-entry byte -> token id table; anything outside the alphabet -> <unk>
    (``set_default_index``, data_processing.py:345)."""
    lut = np.full(256, UNK, dtype=np.int64)
    for i, ch in enumerate(ALL_AMINO_ACIDS):
        lut[ord(ch)] = 4 + i
    return lut


def tokenize(seq: str) -> np.ndarray:
    """SimpleCharacterTokenizer with sos/eos (data_processing.py:50-54,61)."""
    raw = np.frombuffer(seq.encode("latin-1", errors="replace"), dtype=np.uint8)
    ids = char_lut()[raw]
    return np.concatenate([[SOS], ids, [EOS]]).astype(np.int64)


def needs_crop(seq_len_chars: int, max_length: int) -> bool:
    """SentenceRandomCrop draws only when len(ids) > max_length (:78-83)."""
    return seq_len_chars + 2 > max_length


def corrupt_sample(seq: str, ann: np.ndarray, max_length: int, *, crop_start: int,
                   tok_mask: np.ndarray, tok_rand: np.ndarray, coin: float,
                   keep: np.ndarray, add: np.ndarray):
    """One sample -> (X_local, X_global, Y_local, Y_global, W_local, W_global).

    crop_start : the ``randint(0, len-max)`` draw (ignored when no crop happens)
    tok_mask   : the ``multinomial([.95,.05], n)`` draw, 0/1, n = len after crop
    tok_rand   : the ``randint(3, 26, (n,))`` draw
    coin       : the ``rand(1)`` draw; annotations are noised iff coin > 0.5
    keep, add  : the two ``multinomial`` draws of length A (unused when coin <= 0.5)
    """
    ids = tokenize(seq)
    if len(ids) > max_length:                                       # :79-83
        ids = ids[crop_start:crop_start + max_length]
    n = len(ids)
    m = (tok_mask[:n] != 0) & (ids != PAD) & (ids != SOS) & (ids != EOS)   # :101-103
    masked = np.where(m, tok_rand[:n], ids)                         # :104-105
    x_local = np.zeros(max_length, dtype=np.int64)
    y_local = np.zeros(max_length, dtype=np.int64)
    x_local[:n] = masked                                            # :165 right pad with <pad>=0
    y_local[:n] = ids                                               # :167
    ann = np.asarray(ann).astype(np.int64)
    if coin > 0.5:                                                  # :128
        x_global = ((ann + add.astype(np.int64)) * keep.astype(np.int64)).astype(np.float32)  # :139-140
    else:
        x_global = np.zeros(ann.shape, dtype=np.float32)            # :127
    y_global = ann.astype(np.float32)                               # :173,179
    w_local = (y_local != PAD).astype(np.float64)                   # :175
    w_global = np.full(ann.shape, float(ann.any()), dtype=np.float64)   # :176
    return x_local, x_global, y_local, y_global, w_local, w_global


def corrupt_batch(seqs, anns, max_length, draws):
    """Batch version; ``draws`` is a list of per-sample dicts with the keys of
    :func:`corrupt_sample`.  Returns the six stacked arrays (default_collate)."""
    outs = [corrupt_sample(s, a, max_length, **d) for s, a, d in zip(seqs, anns, draws)]
    return tuple(np.stack([o[i] for o in outs]) for i in range(6))
