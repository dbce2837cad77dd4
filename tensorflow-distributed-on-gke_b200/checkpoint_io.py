"""Weight files under the reference's `save_weights(prefix)` / `load_weights(prefix)` call shape
(checkpoint.py:86, __main__.py:87).  Variables are stored under the Keras object-graph keys of
saved_weights/*/model_weights.index (`<name>/.ATTRIBUTES/VARIABLE_VALUE`), in a NumPy .npz container
next to the prefix; the TF tensor-bundle (SSTable) container itself is SURVEY §8(f) rank 1, not built yet."""
from __future__ import annotations

from pathlib import Path

import numpy as np

SUFFIX = "/.ATTRIBUTES/VARIABLE_VALUE"


def save_weights(model, prefix):
    p = Path(str(prefix) + ".b200xf.npz")
    p.parent.mkdir(parents=True, exist_ok=True)
    np.savez(p, **{k + SUFFIX: v for k, v in model.get_weights_dict().items()})


def load_weights(model, prefix):
    p = Path(str(prefix) + ".b200xf.npz")
    if not p.exists():
        raise FileNotFoundError(f"{p} (TF tensor-bundle checkpoints are not readable yet)")
    z = np.load(p)
    model.set_weights_dict({k[:-len(SUFFIX)]: z[k] for k in z.files})
