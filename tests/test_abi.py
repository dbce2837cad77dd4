"""CPU-side checks of the C-ABI boundary: the library loads, exports every symbol include/b200xf.h declares,
and refuses to run without a GPU (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import pytest
import torch

import b200xf

ROOT = Path(__file__).resolve().parent.parent


def _declared():
    text = (ROOT / "include" / "b200xf.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200xf_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(str(b200xf.LIB_PATH))
    names = _declared()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200xf.h but not exported"


def test_ctypes_signatures_cover_the_header():
    assert sorted(b200xf._lib.SIGNATURES) == _declared()


def test_version_and_error_string():
    lib = b200xf._lib.load()
    assert lib.b200xf_version() >= 100
    assert isinstance(lib.b200xf_last_error(), bytes)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu():
    m = b200xf.Transformer(2, 32, 4, 64, 50, 40, 64, 64, 0.0)
    with pytest.raises(b200xf.B200XFError):
        m.trainable_variables
    lib = b200xf._lib.load()
    hp = b200xf._lib.HParams(2, 32, 4, 64, 50, 40, 64, 64, 0.0, 0.9, 0.98, 1e-9, 4000, 256)
    ctx = C.c_void_p()
    assert lib.b200xf_ctx_create(C.byref(hp), 0, None, C.byref(ctx)) != 0
    assert b"no CUDA device" in lib.b200xf_last_error()


def test_host_logic_masks_and_schedule():
    assert b200xf.create_look_ahead_mask(3).tolist() == [[0, 1, 1], [0, 0, 1], [0, 0, 0]]
    m = b200xf.create_padding_mask(torch.tensor([[7, 6, 0, 0, 1]]))
    assert tuple(m.shape) == (1, 1, 1, 5) and m[0, 0, 0].tolist() == [0, 0, 1, 1, 0]
    s = b200xf.TransformerLearningRateSchedule(128)
    assert s(0) == 0.0 and abs(float(s(4000)) - 1.39754254e-03) < 1e-8
    assert b200xf.CustomSchedule is b200xf.TransformerLearningRateSchedule
    t = b200xf.Transformer(num_layers=2, d_model=32, num_heads=4, dff=64, input_vocab_size=50, target_vocab_size=40,
                           pe_input=64, pe_target=64, rate=0.0)
    assert t.hparams["layers_count"] == 2 and t.hparams["droput_rate"] == 0.0
    with pytest.raises(AssertionError):
        b200xf.Transformer(2, 30, 4, 64, 50, 40, 64, 64, 0.0)
    _, la, _ = t.create_masks(torch.tensor([[1, 2, 3]]), torch.tensor([[5, 9, 0]]))
    assert la[0, 0].tolist() == [[0, 1, 1], [0, 0, 1], [0, 0, 1]]
