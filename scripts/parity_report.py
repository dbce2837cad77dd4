"""Prints a parity report of the CUDA train step against the CPU oracle (run on the GPU box)."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200xf
from oracle import transformer_oracle as O

TUT = dict(layers_count=4, embedding_size=128, heads_count=8, feed_forward_hidden_size=512, input_vocabulary_size=7765,
           target_vocabulary_size=7010, maximum_input_length=1000, maximum_target_length=1000, droput_rate=0.0)


def rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


SMALL = dict(layers_count=2, embedding_size=64, heads_count=4, feed_forward_hidden_size=128, input_vocabulary_size=203,
             target_vocabulary_size=157, maximum_input_length=64, maximum_target_length=64, droput_rate=0.0)


def main():
    hp, B, S, T = (SMALL, 4, 20, 17) if "small" in sys.argv else (TUT, 64, 40, 40)
    model = b200xf.Transformer(**hp, seed=0)
    w0 = model.get_weights_dict()
    oracle = O.OracleTransformer(**hp).load(w0)
    oracle16 = O.OracleTransformer(**hp, emulate_bf16=True).load(w0)
    inp, tgt = b200xf.synthetic_batch(B, S, T + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=77, ragged=True)
    out = model.forward_backward_device(inp, tgt, 1).cpu().numpy()
    grads = model.get_slot_dict("grad")
    loss_ref, logits_ref, gref = O.train_step(oracle, inp, tgt, 1)
    loss16, _, g16 = O.train_step(oracle16, inp, tgt, 1)
    print(f"loss cuda {out[0]:.6f} oracle {loss_ref:.6f} oracle-bf16emu {loss16:.6f}")
    rows = []
    for n in gref:
        rows.append((rel(grads[n], gref[n]), rel(g16[n], gref[n]), float(np.linalg.norm(gref[n])), n))
    rows.sort(reverse=True)
    print("grad rel-L2: cuda-vs-fp32 | bf16emu-vs-fp32 | ||g_ref|| | name")
    for r in rows[:25]:
        print(f"  {r[0]:.4f}  {r[1]:.4f}  {r[2]:.3e}  {r[3]}")
    print("  median", np.median([r[0] for r in rows]), "emu median", np.median([r[1] for r in rows]))
    allg = np.concatenate([grads[n].ravel() for n in gref]); allr = np.concatenate([gref[n].ravel() for n in gref]); all16 = np.concatenate([g16[n].ravel() for n in gref])
    print("global grad rel-L2 cuda", rel(allg, allr), "bf16emu", rel(all16, allr))
    # multi-step delta-w
    model = b200xf.Transformer(**hp, seed=0); oracle = O.OracleTransformer(**hp).load(w0); oracle16 = O.OracleTransformer(**hp, emulate_bf16=True).load(w0)
    for m in (model, oracle, oracle16):
        m.iterations = 2000
    for step in range(4):
        inp, tgt = b200xf.synthetic_batch(B, S, T + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=1235 + step, ragged=True)
        r = model.train_step_device(inp, tgt, 1).cpu().numpy()
        lr_, _, gref = O.train_step(oracle, inp, tgt, 1)
        l16, _, _ = O.train_step(oracle16, inp, tgt, 1)
        print(f"step {step}: loss cuda {r[0]:.5f} oracle {lr_:.5f} emu {l16:.5f}")
    w = model.get_weights_dict()
    for thr in (1e-7, 1e-6, 1e-5):
        for label, other in (("cuda", w), ("bf16emu", {k: p.detach().numpy() for k, p in oracle16.params.items()})):
            num = den = agree = tot = 0.0
            for name, p in oracle.params.items():
                ref = p.detach().numpy(); d_ref = ref - w0[name]; d_got = other[name] - w0[name]
                sel = np.abs(gref[name]) > thr
                if sel.any():
                    num += float(np.sum((d_got[sel] - d_ref[sel]) ** 2)); den += float(np.sum(d_ref[sel] ** 2))
                    wgt = np.abs(gref[name][sel]); agree += float(np.sum(wgt * (np.sign(d_got[sel]) == np.sign(d_ref[sel])))); tot += float(np.sum(wgt))
            print(f"  |g_ref|>{thr:g} {label}: delta-w rel-L2 {np.sqrt(num / den):.4f}, weighted sign agreement {agree / tot:.4f}")


if __name__ == "__main__":
    main()
