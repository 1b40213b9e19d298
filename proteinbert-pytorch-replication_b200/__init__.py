"""proteinbert-pytorch-replication_b200: ProteinBERT pretraining forward/backward on B200 (sm_100a).

The directory name contains hyphens, so import it through the shim module at the repository root:

    import proteinbert_b200 as pb
    model = pb.ProteinBERT(sequences_length=512, num_annotations=8943, local_dim=128, global_dim=512,
                           key_dim=64, num_heads=4, num_blocks=6, precision="bf16")

Public surface = the reference's ``ProteinBERT/modules.py`` classes, the loss expression of
``utils.py:293-294`` (``pretraining_loss``), the corruption/padding of ``data_processing.py`` and a
data-parallel gradient reducer.  All arithmetic runs in libpb_b200.so (C ABI: include/pb_b200.h).
"""
from . import _lib, ops
from ._lib import PBError, launch_count
from .modules import GlobalAttention, GlobalAttentionHead, ProteinBERT, ProteinBERTBlock
from .training import FusedAdam, GradAllReducer, GraphedPretrainStep, WarmupThenPlateau, pretrain, pretraining_loss, pretraining_step
from . import data_processing

__all__ = ["ProteinBERT", "ProteinBERTBlock", "GlobalAttention", "GlobalAttentionHead", "pretraining_loss",
           "pretraining_step", "pretrain", "WarmupThenPlateau", "GraphedPretrainStep", "GradAllReducer", "FusedAdam", "data_processing", "ops", "PBError", "launch_count"]
