"""ctypes binding of libb200xf.so (C ABI declared in include/b200xf.h).

There is deliberately no fallback: if the shared library is missing or no CUDA device is
present, calls fail loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("B200XF_LIB", _HERE / "libb200xf.so"))


class B200XFError(RuntimeError):
    pass


class HParams(C.Structure):
    _fields_ = [
        ("layers_count", C.c_int32), ("embedding_size", C.c_int32), ("heads_count", C.c_int32),
        ("feed_forward_hidden_size", C.c_int32), ("input_vocabulary_size", C.c_int32),
        ("target_vocabulary_size", C.c_int32), ("maximum_input_length", C.c_int32),
        ("maximum_target_length", C.c_int32), ("droput_rate", C.c_float), ("adam_beta_1", C.c_float),
        ("adam_beta_2", C.c_float), ("adam_epsilon", C.c_float), ("warmup_steps", C.c_int32),
        ("max_tokens", C.c_int32),
    ]


_P = C.c_void_p
_I = C.c_int
_F = C.c_float
_L = C.c_int64
_SZ = C.c_size_t

# name -> (restype, argtypes); mirrors include/b200xf.h one to one.
SIGNATURES = {
    "b200xf_last_error": (C.c_char_p, []),
    "b200xf_version": (_I, []),
    "b200xf_launch_count": (C.c_longlong, []),
    "b200xf_ctx_create": (_I, [C.POINTER(HParams), _I, _P, C.POINTER(_P)]),
    "b200xf_ctx_destroy": (_I, [_P]),
    "b200xf_device_bytes": (_I, [C.POINTER(HParams), C.POINTER(_L), C.POINTER(_L)]),
    "b200xf_param_count": (_I, [_P]),
    "b200xf_param_info": (_I, [_P, _I, C.c_char_p, _SZ, C.POINTER(_L), C.POINTER(_L), C.POINTER(_L), C.POINTER(_L), C.POINTER(C.c_int32)]),
    "b200xf_param_total": (_L, [_P]),
    "b200xf_param_ptr": (_P, [_P]),
    "b200xf_grad_ptr": (_P, [_P]),
    "b200xf_adam_m_ptr": (_P, [_P]),
    "b200xf_adam_v_ptr": (_P, [_P]),
    "b200xf_params_changed": (_I, [_P]),
    "b200xf_set_iterations": (_I, [_P, _L]),
    "b200xf_get_iterations": (_I, [_P, C.POINTER(_L)]),
    "b200xf_forward": (_I, [_P, _P, _P, _I, _I, _I, _I, _P]),
    "b200xf_decode_begin": (_I, [_P, _P, _I, _I, _I]),
    "b200xf_decode_step": (_I, [_P, _P, _I, _P, _P]),
    "b200xf_attention_weights": (_I, [_P, _I, _I, _P]),
    "b200xf_argmax_rows": (_I, [_P, _I, _I, _I, _P, _P]),
    "b200xf_train_step_device": (_I, [_P, _P, _P, _I, _I, _I, _I, _P]),
    "b200xf_train_step_host": (_I, [_P, _P, _P, _I, _I, _I, _I, _P]),
    "b200xf_eval_step_device": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "b200xf_forward_backward_device": (_I, [_P, _P, _P, _I, _I, _I, _I, _P]),
    "b200xf_synchronize": (_I, [_P]),
    "b200xf_set_dropout_seed": (_I, [_P, C.c_uint64]),
    "b200xf_dropout_mask": (_I, [_P, _I, _I, _I, _P]),
    "b200xf_comm_unique_id": (_I, [_P]),
    "b200xf_comm_init": (_I, [_P, _P, _I, _I]),
    "b200xf_comm_destroy": (_I, [_P]),
    "b200xf_use_external_state": (_I, [_P, _P, _P]),
    "b200xf_nvls_attach": (_I, [_P, _P, _P, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), _I, _I, _I]),
    "b200xf_nvls_error": (_I, [_P]),
    "b200xf_debug_nvls_update": (_I, [_P, _L, _L, _I, _I]),
    "b200xf_gather_optimizer_state": (_I, [_P]),
    "b200xf_optimizer_state_complete": (_I, [_P]),
    "b200xf_allreduce_bucket": (_I, [_P, _P, _L]),
    "b200xf_gemm_bf16": (_I, [_P, _I, _I, _P, _I, _I, _P, _I, _I, _I, _I, _I, _I, _P, _I, _P, _I, _I, _F, _I, _I, _P]),
    "b200xf_dense_layernorm_fwd": (_I, [_P, _I, _P, _I, _P, _P, _I, _P, _I, _P, _P, _P, _I, _P, _P, _I, _I, _I, _F, _P]),
    "b200xf_dense_wgrad_bias": (_I, [_P, _I, _P, _I, _P, _I, _P, _P, _SZ, _I, _I, _I, _P]),
    "b200xf_assemble_padded": (_I, [_P, _P, _I, _I, _I, _P, _P, _P]),
    "b200xf_debug_gemm_trace": (_I, [_P]),
    "b200xf_debug_gemm_trace_ring": (_I, [_P, _I]),
    "b200xf_gemm_set_pair_mode": (_I, [_I]),
    "b200xf_attention_fwd": (_I, [_P, _I, _P, _I, _P, _I, _P, _I, _P, _P, _I, _I, _I, _I, _I, _I, _P]),
    "b200xf_attention_bwd": (_I, [_P, _I, _P, _I, _P, _I, _P, _I, _P, _P, _I, _I, _I, _I, _I, _I, _P, _I, _P, _I, _P, _I, _P, _I, _P, _P, _P]),
    "b200xf_layernorm_residual_fwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _F, _P]),
    "b200xf_layernorm_bwd_scratch_bytes": (_SZ, [_I, _I]),
    "b200xf_layernorm_residual_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P]),
    "b200xf_layernorm_residual_bwd_noalloc": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P]),
    "b200xf_embed_pe_fwd": (_I, [_P, _P, _I, _P, _P, _I, _I, _I, _P]),
    "b200xf_embed_bwd_scratch_bytes": (_SZ, [_I]),
    "b200xf_embed_pe_bwd": (_I, [_P, _P, _P, _I, _I, _I, _P, _SZ, _P]),
    "b200xf_vocab_ce_fwd_bwd": (_I, [_P, _I, _P, _I, _I, _I, _I, _P, _P, _P]),
    "b200xf_vocab_ce_f32": (_I, [_P, _I, _P, _I, _I, _I, _P, _P, _P]),
    "b200xf_adam_step": (_I, [_P, _P, _P, _P, _P, _L, _P, _I, _I, _F, _F, _F, _I, _P]),
    "b200xf_crc32c": (C.c_uint32, [_P, _SZ, C.c_uint32]),
    "b200xf_positional_encoding_host": (_I, [_P, _I, _I]),
}

_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise B200XFError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            f"(or `make -C {(_HERE / 'csrc')}`).  There is no CPU fallback.")
    lib = C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the ABI and the header disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int):
    if rc != 0:
        raise B200XFError(load().b200xf_last_error().decode("utf-8", "replace"))


def ptr(t):
    """Device (or host) address of a torch tensor / None."""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def stream_ptr(stream=None):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)
