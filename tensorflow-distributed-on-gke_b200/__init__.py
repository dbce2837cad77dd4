"""B200-native drop-in for the data-parallel Transformer train step of ddrakard/tensorflow-distributed-on-gke.

Python host code mirroring the reference's `transformer_model.py` / `__main__.py` surface over a C-ABI
CUDA library (libb200xf.so, sm_100a).  No CPU fallback: without the built library and a GPU, calls raise.
"""
from . import _lib
from ._lib import B200XFError, LIB_PATH
from .batching import BatchAssembler, pad_to_longest, padded_length, ragged_to_flat, replica_rows
from .cluster import B200Cluster, launch
from .synthetic import CONFIGS, step_flops, synthetic_batch
from .training import (Adam, B200Strategy, CustomSchedule, Mean, ReduceOp, StepResult, TransformerLearningRateSchedule,
                       make_distributed_test_step, make_distributed_train_step, make_test_step, make_train_step)
from .transformer_model import (AttentionWeights, Transformer, accuracy_function, build_model, greedy_decode, greedy_decode_batch, create_look_ahead_mask, create_padding_mask,
                                local_loss_function, loss_function, positional_encoding)

__all__ = [
    "Transformer", "AttentionWeights", "greedy_decode", "greedy_decode_batch", "build_model", "local_loss_function", "loss_function", "accuracy_function", "positional_encoding",
    "create_padding_mask", "create_look_ahead_mask", "TransformerLearningRateSchedule", "CustomSchedule", "Adam",
    "B200Strategy", "ReduceOp", "Mean", "StepResult", "make_train_step", "make_distributed_train_step", "make_distributed_test_step", "make_test_step",
    "synthetic_batch", "step_flops", "CONFIGS", "B200XFError", "LIB_PATH",
    "BatchAssembler", "pad_to_longest", "padded_length", "ragged_to_flat", "replica_rows", "B200Cluster", "launch",
]
