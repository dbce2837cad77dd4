#!/usr/bin/env python
"""Benchmark of the data-parallel Transformer train step (BASELINE.json metric: target tokens/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config base|tutorial|big|long] [--impl reference]

One process per GPU (torchrun sets RANK/LOCAL_RANK/WORLD_SIZE for N > 1).  A "step" is one train_step
(forward + masked loss + backward + gradient all-reduce + Adam) over one synthetic batch.  Prints ONE JSON line.
  value    device-timed whole-job target tokens/s, token batches already resident in HBM
  e2e      the same metric through the HOST-buffer C-ABI call (H2D of the tokens and D2H of the loss inside the timed region)
  roofline the tcgen05 GEMM kernel: algorithmic FLOPs of the step's GEMM launches (with the step's own epilogues) / their CUDA-event
           time in an isolated replay vs the measured BURST bf16 peak; `in_step`: the same launches' durations inside a replayed
           train step (device time stamps) vs the SUSTAINED peak; `secondary`: the HBM-bound kernels against the measured copy rate
  parity   computed OUTSIDE the timed region on the measured configuration: the step's loss and logits against the CPU oracle on
           this rank's batch, one Adam step's weight change against the oracle's (1 GPU), and for N > 1 that every rank holds
           bit-identical weights and that the reduced loss equals the sum of the per-rank oracle losses
  allreduce (N > 1) gradient bytes per step, NCCL bus bandwidth measured on the bucket sizes, step time with and without collectives
  cpu_baseline  the CPU oracle (a torch fp32 restatement of the reference; TensorFlow itself is not installable here) on the host cores
`--impl reference` times that CPU restatement alone, same metric/config, the workload's full per-GPU batch per step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "target tokens/sec, Transformer train_step"
UNIT = "tokens/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--config", default="base", choices=["tutorial", "base", "big", "long"])
    ap.add_argument("--impl", default="b200xf", choices=["b200xf", "reference"])
    ap.add_argument("--batch", type=int, default=0, help="override sentences per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--hbm-gb", type=float, default=150.0, help="long configuration: device memory the batch is sized to (BASELINE.json configs[4]: ~150 GB)")
    ap.add_argument("--no-allreduce-probe", action="store_true")
    ap.add_argument("--ragged", action="store_true", help="padded batches (lengths U{L/2..L}) instead of full-length ones")
    return ap.parse_args()


def config_dict(name, cfg, B, n_gpus, ragged):
    return {"workload": f"Transformer-{name} train_step: {cfg['layers_count']}+{cfg['layers_count']} layers, d_model {cfg['embedding_size']}, "
                        f"dff {cfg['feed_forward_hidden_size']}, {cfg['heads_count']} heads, vocab {cfg['input_vocabulary_size']}/{cfg['target_vocabulary_size']}, "
                        f"seq {cfg['S']}, {B * cfg['T']} target tokens/GPU, dropout 0",
            "baseline_config": "BASELINE.json configs[2]" if name == "base" else name,
            "global_batch_tokens": B * cfg["T"] * n_gpus, "batch_per_gpu": B, "seq_len": cfg["S"],
            "parallelism": f"dp{n_gpus}", "padding": "ragged" if ragged else "none",
            "l2": "per-step working set (weights+grads+Adam+activations, >1 GB at base) exceeds the 126 MB L2; no explicit flush"}


# ---------------------------------------------------------------------------------------------- CPU reference arm
def _oracle_hp(cfg):
    return dict(layers_count=cfg["layers_count"], embedding_size=cfg["embedding_size"], heads_count=cfg["heads_count"],
                feed_forward_hidden_size=cfg["feed_forward_hidden_size"], input_vocabulary_size=cfg["input_vocabulary_size"],
                target_vocabulary_size=cfg["target_vocabulary_size"], maximum_input_length=max(1000, cfg["S"]),
                maximum_target_length=max(1000, cfg["T"] + 1), droput_rate=0.0)


def cpu_reference(name, cfg, steps, warmup, sample_sentences, threads=None, ragged=False):
    """The reference's train_step restated on the CPU (oracle/), on `sample_sentences` sentences of the workload per step."""
    import statistics
    import torch
    from oracle import transformer_oracle as O
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    hp = _oracle_hp(cfg)
    model = O.OracleTransformer(**hp).load(O.init_parameters(hp["layers_count"], hp["embedding_size"], hp["feed_forward_hidden_size"],
                                                             hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=0))
    model.iterations = 1
    Bs = sample_sentences
    times = []
    for i in range(warmup + steps):
        inp, tgt = O.synthetic_batch(Bs, cfg["S"], cfg["T"] + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=1234 + i, ragged=ragged)
        t0 = time.perf_counter()
        O.train_step(model, inp, tgt, 1)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    per_step = statistics.median(times)
    whole = "the workload's full per-GPU batch" if Bs == cfg["B"] else f"{Bs} of the workload's {cfg['B']} sentences"
    return dict(value=Bs * cfg["T"] / per_step, unit=UNIT, cores=cores, kind="port",
                sample=f"{whole} per step ({Bs * cfg['T']} target tokens), {warmup} warm-up + {len(times)} timed steps (median), "
                       f"torch-CPU fp32 restatement of the reference (TensorFlow not installable here), {cores} threads"), per_step


REF_SAMPLE = {"tutorial": 64, "base": 32, "big": 4, "long": 1}   # sentences per CPU step (tutorial / base: the whole per-GPU batch)


def cpu_baseline_block(name, cfg, quick):
    """cpu_baseline of the JSON line: the workload itself on all host threads, plus the tutorial configuration exactly as
    BASELINE.md section 2 specifies it (64 x 40 ragged batch, 3 warm-up + 20 timed steps, all threads and one thread)."""
    cb, _ = cpu_reference(name, cfg, 2 if quick else 3, 1, REF_SAMPLE[name])
    import b200xf
    tut = dict(b200xf.CONFIGS["tutorial"])
    allt, _ = cpu_reference("tutorial", tut, 5 if quick else 20, 3, 64, ragged=True)
    one, _ = cpu_reference("tutorial", tut, 3 if quick else 10, 2, 64, threads=1, ragged=True)
    cb["tutorial_config"] = {"all_threads": {"value": allt["value"], "cores": allt["cores"], "sample": allt["sample"]},
                             "one_thread": {"value": one["value"], "cores": 1, "sample": one["sample"]}, "unit": UNIT,
                             "what": "BASELINE.md section 2: reference tutorial Transformer, input [64,40] / target [64,41], lengths U{20..40}"}
    return cb


def run_reference(args):
    import b200xf  # only for the config table (pure Python import; no GPU work)
    cfg = dict(b200xf.CONFIGS[args.config])
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    sample = REF_SAMPLE[args.config]
    steps = max(1, min(args.steps, 20 if args.config == "tutorial" else 3)); warmup = max(1, min(args.warmup, 3 if args.config == "tutorial" else 1))
    cb, per_step = cpu_reference(args.config, cfg, steps, warmup, sample)
    B = args.batch or cfg["B"]
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
            "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": config_dict(args.config, cfg, B, args.gpus, False), "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows = []; self.proc = None; self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def window(self, t0, t1):
        self.t0, self.t1 = t0, t1

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try: self.proc.wait(timeout=2)
        except Exception: self.proc.kill()
        t0, t1 = getattr(self, "t0", 0.0), getattr(self, "t1", float("inf"))
        inside = [r for t, r in self.rows if t0 <= t <= t1]
        self.rows = inside if len(inside) >= 2 else [r for _, r in self.rows]  # very short runs: fall back to every sample under load
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------------------------- roofline legs
def gemm_shapes(cfg, Ns, Nt):
    """Every GEMM launch of one train step as (M, N, K, layout, epilogue): forward y = xW + b (ReLU for FFN1, + residual for the
    out-projections and FFN2), dgrad dx = dy W^T (+ residual gradient, or the ReLU mask for FFN2's), wgrad dW += x^T dy (fp32)."""
    L, d, dff, V = cfg["layers_count"], cfg["embedding_size"], cfg["feed_forward_hidden_size"], cfg["target_vocabulary_size"]
    dense = []  # (tokens, K_in, N_out, forward epilogue, dgrad epilogue)
    for _ in range(L):
        dense += [(Ns, d, 3 * d, "bias", "resid"), (Ns, d, d, "bias+resid", "none"), (Ns, d, dff, "bias+relu", "resid"), (Ns, dff, d, "bias+resid", "relumask")]
    dense += [(Ns, d, L * 2 * d, "bias", "none")]
    for _ in range(L):
        dense += [(Nt, d, 3 * d, "bias", "resid"), (Nt, d, d, "bias+resid", "none"), (Nt, d, d, "bias", "resid"), (Nt, d, d, "bias+resid", "none"),
                  (Nt, d, dff, "bias+relu", "resid"), (Nt, dff, d, "bias+resid", "relumask")]
    dense += [(Nt, d, V, "bias", "none")]
    shapes = []
    for tok, K, N, fe, de in dense:
        shapes += [(tok, N, K, "fwd", fe), (tok, K, N, "dgrad", de), (K, N, tok, "wgrad", "accumulate")]
    return shapes


def gemm_roofline(cfg, Ns, Nt, reps=3):
    import torch
    from b200xf import kernels as Kn
    shapes = gemm_shapes(cfg, Ns, Nt)
    bufs = {}
    LAY = {"fwd": (False, True), "dgrad": (False, False), "wgrad": (True, True)}
    pad = lambda n: (n + 7) // 8 * 8

    def operands(M, N, K, lay, epi):
        key = (M, N, K, lay, epi)
        if key not in bufs:
            a_mn, b_mn = LAY[lay]
            a = torch.randn((K, pad(M)) if a_mn else (M, pad(K)), device="cuda").to(torch.bfloat16)
            b = torch.randn((K, pad(N)) if b_mn else (N, pad(K)), device="cuda").to(torch.bfloat16)
            a = a[:, :M] if a_mn else a[:, :K]; b = b[:, :N] if b_mn else b[:, :K]
            out = torch.zeros((M, pad(N)), device="cuda", dtype=torch.float32 if lay == "wgrad" else torch.bfloat16)[:, :N]
            bias = torch.randn(pad(N), device="cuda") if "bias" in epi else None
            resid = torch.randn((M, pad(N)), device="cuda").to(torch.bfloat16)[:, :N] if ("resid" in epi or epi == "relumask") else None
            bufs[key] = (a, b, out, a_mn, b_mn, bias, resid)
        return bufs[key]

    def run_all():
        for M, N, K, lay, epi in shapes:
            a, b, out, a_mn, b_mn, bias, resid = operands(M, N, K, lay, epi)
            Kn.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, accumulate=(lay == "wgrad"), bias=bias, relu=("relu" in epi and epi != "relumask"),
                    resid=resid, resid_op=2 if epi == "relumask" else 1)
    run_all(); torch.cuda.synchronize()
    # replay through a CUDA graph so that the Python/ctypes call cost does not leave the GPU idle between launches
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        run_all(); side.synchronize()
        with torch.cuda.graph(graph, stream=side):
            run_all()
        graph.replay(); side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(reps):
            graph.replay()
        e1.record(side); side.synchronize()
    ms = e0.elapsed_time(e1) / reps
    flops = sum(2.0 * M * N * K for M, N, K, _, _ in shapes)
    return flops, ms, len(shapes)


def gemm_in_step(hp, B, S, T, max_tokens):
    """Durations of the step's own GEMM launches inside a graph-replayed train step: every GEMM CTA stamps %globaltimer
    (b200xf_debug_gemm_trace_ring; the stamps are part of the captured graph).  A launch's duration = last CTA exit - the
    median CTA's dependency-resolved stamp.  Diagnostic: stamping adds a little time; the step is not timed here."""
    import torch
    import b200xf
    from b200xf._lib import load, ptr
    lib = load()
    SLOTS = 512
    ring = torch.zeros(SLOTS * 148 * 8, device="cuda", dtype=torch.int64)
    lib.b200xf_debug_gemm_trace_ring(ptr(ring), SLOTS)
    try:
        model = b200xf.Transformer(**hp, max_tokens=max_tokens, seed=0)
        model.trainable_variables
        model.iterations = 1
        inp, tgt = b200xf.synthetic_batch(B, S, T + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=1, ragged=False)
        inp, tgt = torch.from_numpy(inp).cuda(), torch.from_numpy(tgt).cuda()
        for _ in range(6):
            model.train_step_device(inp, tgt)
        torch.cuda.synchronize(); ring.zero_(); torch.cuda.synchronize()
        model.train_step_device(inp, tgt); torch.cuda.synchronize()
    finally:
        lib.b200xf_debug_gemm_trace_ring(None, 0)
    t = ring.cpu().view(SLOTS, 148, 8)
    flops = busy_ns = 0.0; n = 0; first = None; last = 0
    for s in range(SLOTS):
        used = t[s, :, 0] > 0
        if not bool(used.any()):
            continue
        u = t[s][used]; d = int(t[s, 0, 7])
        M, N, K = d >> 44, (d >> 24) & 0xFFFFF, (d >> 4) & 0xFFFFF
        flops += 2.0 * M * N * K; busy_ns += float(u[:, 6].max() - u[:, 1].median()); n += 1
        first = int(u[:, 0].min()) if first is None else min(first, int(u[:, 0].min())); last = max(last, int(u[:, 6].max()))
    model.close()
    return flops, busy_ns / 1e6, n, (last - (first or 0)) / 1e6


def secondary_rooflines(cfg, Nt, n_params, peak_hbm):
    """The HBM-bound kernels, each timed alone (CUDA graph of launches over rotating buffer sets larger than L2, CUDA events):
    algorithmic bytes per launch / duration against the measured copy rate."""
    import torch
    from b200xf import kernels as Kn
    d, V = cfg["embedding_size"], cfg["target_vocabulary_size"]
    out = []
    traffic = {}
    tp = ROOT / "profiles" / "r02_traffic.json"   # dram__bytes_read.sum + dram__bytes_write.sum per launch from ncu --set full captures
    if tp.exists():
        traffic = json.loads(tp.read_text())

    def timed(fn, sets, reps=4):
        st = torch.cuda.Stream(); g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(st):
            for i in range(sets): fn(i)
            st.synchronize()
            with torch.cuda.graph(g, stream=st):
                for i in range(sets): fn(i)
            g.replay(); st.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(reps): g.replay()
            e1.record(st); st.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (reps * sets)   # us per launch

    def entry(kernel, bytes_alg, us, how):
        ach = bytes_alg / (us * 1e-6) / 1e9
        out.append({"kernel": kernel, "bound": "hbm", "achieved": ach, "peak": peak_hbm, "unit": "GB/s", "frac": ach / peak_hbm,
                    "traffic": traffic.get(kernel), "bytes_per_launch": bytes_alg, "avg_launch_us": us, "how": how})

    rows = Nt
    sets = max(2, int(300e6 // (rows * d * 2 * 3)) + 1)   # x, resid, y of all sets together exceed the 126 MB L2
    g = torch.Generator(device="cuda").manual_seed(1)
    xs = [torch.randn(rows, d, device="cuda", generator=g).to(torch.bfloat16) for _ in range(sets)]
    rs = [torch.randn(rows, d, device="cuda", generator=g).to(torch.bfloat16) for _ in range(sets)]
    gamma = torch.ones(d, device="cuda"); beta = torch.zeros(d, device="cuda")
    lib = __import__("b200xf")._lib.load(); from b200xf._lib import ptr, stream_ptr, check
    ys = [torch.empty_like(x) for x in xs]; mean = torch.empty(rows, device="cuda"); rstd = torch.empty(rows, device="cuda")
    us = timed(lambda i: check(lib.b200xf_layernorm_residual_fwd(ptr(xs[i]), None, ptr(gamma), ptr(beta), ptr(ys[i]), None, ptr(mean), ptr(rstd), rows, d, 1e-6, stream_ptr())), sets)
    entry("ln_fwd_kernel", rows * d * 2 * 2, us, f"{rows}x{d} bf16, read z + write y, {sets} rotating buffer sets")
    scratch = Kn.layernorm_bwd_scratch(rows, d, "cuda")
    dzs = [torch.empty_like(x) for x in xs]; dg = torch.empty(d, device="cuda"); db = torch.empty(d, device="cuda"); dc = torch.empty(d, device="cuda")
    us = timed(lambda i: check(lib.b200xf_layernorm_residual_bwd(ptr(xs[i]), ptr(rs[i]), ptr(mean), ptr(rstd), ptr(gamma), ptr(dzs[i]), ptr(dg), ptr(db), ptr(dc), rows, d, ptr(scratch), stream_ptr())), sets)
    entry("ln_bwd_kernel", rows * d * 2 * 3, us, f"{rows}x{d} bf16, read dy + z, write dz (+ partial-sum reduce launch), {sets} rotating buffer sets")
    del xs, rs, ys, dzs
    # cross-entropy: logits [rows, V] bf16 read once and overwritten with the gradient (262 MB at base: twice the L2)
    ld = (V + 7) // 8 * 8
    lg = [torch.randn(rows, ld, device="cuda", generator=g).to(torch.bfloat16) for _ in range(2)]
    labels = torch.randint(1, V, (rows,), device="cuda", generator=g)
    sc = torch.empty(2 * rows + 64, device="cuda"); o3 = torch.empty(3, device="cuda")
    us = timed(lambda i: check(lib.b200xf_vocab_ce_fwd_bwd(ptr(lg[i]), ld, ptr(labels), rows, V, 1, 1, ptr(o3), ptr(sc), stream_ptr())), 2)
    entry("vocab_ce_kernel", rows * V * 2 * 2, us, f"{rows}x{V} bf16 logits read + gradient written in place (includes the tiny count / finalize launches)")
    del lg
    # Adam over the model's parameter count: w, g, m, v read; w, m, v, bf16 copy, cleared g written = 34 B / parameter
    n = (n_params + 3) // 4 * 4
    w, gr, m, v = (torch.randn(n, device="cuda", generator=g) * 0.01 for _ in range(4)); v.abs_()
    wb = torch.empty(n, device="cuda", dtype=torch.bfloat16); it = torch.tensor([10], device="cuda", dtype=torch.int64)
    us = timed(lambda i: check(lib.b200xf_adam_step(ptr(w), ptr(gr), ptr(m), ptr(v), ptr(wb), n, ptr(it), d, 4000, 0.9, 0.98, 1e-9, 1, stream_ptr())), 1, reps=3)
    entry("adam_kernel", n * 34, us, f"{n} parameters x 34 B in one launch")
    return out


def measured_peaks():
    """(burst bf16 TFLOP/s, sustained bf16 TFLOP/s, HBM GB/s, source)"""
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        j = json.loads(p.read_text())
        burst = float(j.get("bf16_tflops", j.get("bf16_tflops_sustained")))
        return burst, float(j.get("bf16_tflops_sustained", burst)), float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 1590.0, 1400.0, 6650.0, "of fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------- parity (outside the timed region)
PARITY_SENTENCES = {"tutorial": 64, "base": 32, "big": 4, "long": 1}   # sentences of this rank's batch handed to the CPU oracle


def parity_block(model, name, cfg, hp, host_batch, world, rank, dist):
    """Checks of the MEASURED configuration against the CPU oracle, with the weights the timed steps have produced.
    Tolerances are those of tests/test_train_step_gpu.py.  Every rank takes part (its own batch)."""
    import numpy as np
    import torch
    from oracle import transformer_oracle as O
    torch.set_num_threads(max(1, (os.cpu_count() or 1) // max(1, world)))
    out = {"tolerances": {"loss_abs": 1e-2, "logits_rel_l2": 2e-2, "delta_w_rel_l2": 0.15, "cross_rank_weights": "bit-identical"}}
    n = min(PARITY_SENTENCES[name], host_batch[0].shape[0])
    inp, tgt = host_batch[0][:n], host_batch[1][:n]
    # (1) every replica holds the same weights after the timed steps (mirrored variables, cluster/cluster.py:66)
    flat = model._flat("param")
    chk = torch.stack([flat.double().sum(), flat.view(torch.int32).to(torch.int64).sum().double()])   # value sum, bit-pattern sum
    if world > 1:
        allc = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(allc, chk)
        allc = torch.stack(allc)
        out["cross_rank_weight_checksum_max_abs_diff"] = float((allc - allc[0]).abs().max().item())
        out["cross_rank_weights_identical"] = bool((allc == allc[0]).all().item())
    # (2) the oracle takes over the GPU's current state (after a collective gather of the sharded Adam slots)
    model.gather_optimizer_state()
    weights = model.get_weights_dict()
    it0 = model.iterations
    oracle = O.OracleTransformer(**hp).load(weights)
    oracle.iterations = it0
    full_step = (world == 1 and name in ("tutorial", "base"))
    if full_step:
        m, v = model.get_slot_dict("adam_m"), model.get_slot_dict("adam_v")
        for k in oracle.m:
            oracle.m[k] = m[k].copy(); oracle.v[k] = v[k].copy()
    # forward: logits and loss on the same tokens
    logits, _ = model([inp, tgt[:, :-1]], training=False)
    ref_logits, _ = oracle([torch.tensor(inp), torch.tensor(tgt[:, :-1])])
    got = logits.cpu().numpy().astype(np.float64); ref = ref_logits.detach().numpy().astype(np.float64)
    out["logits_rel_l2"] = float(np.linalg.norm(got - ref) / np.linalg.norm(ref))
    # one more (untimed) train step on these tokens: its loss is computed with the weights the oracle was just given
    res = model.train_step_device(inp, tgt, workers_count=world).cpu().numpy()
    if full_step:
        ref_loss, _, gref = O.train_step(oracle, inp, tgt, world)
        w1 = model.get_weights_dict()
        num = den = agree = tot = 0.0
        for k, p_ in oracle.params.items():
            d_ref = p_.detach().numpy() - weights[k]; d_got = w1[k] - weights[k]
            sel = np.abs(gref[k]) > 1e-7
            if sel.any():
                num += float(np.sum((d_got[sel] - d_ref[sel]) ** 2)); den += float(np.sum(d_ref[sel] ** 2))
                wgt = np.abs(gref[k][sel]); agree += float(np.sum(wgt * (np.sign(d_got[sel]) == np.sign(d_ref[sel])))); tot += float(np.sum(wgt))
        out["one_step_delta_w_rel_l2"] = float(np.sqrt(num / max(den, 1e-300)))
        out["one_step_delta_w_sign_agreement"] = float(agree / max(tot, 1e-300))
    else:
        ref_loss = float(O.local_loss_function(torch.tensor(tgt[:, 1:]), ref_logits.detach(), world))
    out["sentences_checked"] = int(n); out["iterations_at_check"] = int(it0)
    out["loss_gpu"] = float(res[0]); out["loss_oracle"] = float(ref_loss); out["loss_abs_diff"] = abs(float(res[0]) - float(ref_loss))
    ok = out["loss_abs_diff"] < 1e-2 and out["logits_rel_l2"] < 2e-2 and out.get("one_step_delta_w_rel_l2", 0.0) < 0.15
    if world > 1:
        # strategy.reduce(SUM, per-replica losses) (__main__.py:131-132) against the sum of the per-rank oracle losses
        t = torch.tensor([float(ref_loss)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t)
        out["sum_loss_gpu"] = float(res[3]); out["sum_loss_oracle"] = float(t.item())
        out["sum_loss_abs_diff"] = abs(out["sum_loss_gpu"] - out["sum_loss_oracle"])
        ok = ok and out["sum_loss_abs_diff"] < 1e-2 and out["cross_rank_weights_identical"]
        flag = torch.tensor([1.0 if ok else 0.0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(flag.item() > 0.5)
    if getattr(model, "_symmetric", None) is not None:   # fused NVLS update: no replica ever timed out at a cross-replica flag
        from b200xf import _lib
        out["nvls_error"] = int(_lib.load().b200xf_nvls_error(model._ctx))
        ok = ok and out["nvls_error"] == 0
    out["ok"] = bool(ok)
    return out


def allreduce_block(model, hp, B, S, T, world, rank, dist, ms_per_step, n_params, dev_batches, steps, solo_fits=True):
    """What the gradient exchange costs (N > 1): bytes, NCCL bus bandwidth on the library's communicator at the bucket
    sizes, and the step with and without collectives (a second, communicator-less model on the same GPU)."""
    import torch
    import b200xf
    from b200xf._lib import load, ptr, check
    lib = load()
    out = {"wire_dtype": "fp32", "grad_bytes_per_step": int(n_params * 4), "bus_bytes_per_step": n_params * 4 * 2.0 * (world - 1) / world,
           "nvlink_peak_gbs_per_direction": 900.0,
           "collective": ("fused NVLS kernel per >= 32 MB bucket: multimem.ld_reduce of the rank's shard + Adam + multimem.st of the new weights to every replica"
                          if getattr(model, "_symmetric", None) is not None else "reduce-scatter + all-gather per >= 32 MB bucket (sharded Adam between them), NCCL")}
    sweep = []
    for mb in (8, 32, 128, 373):
        n = min(mb << 18, (n_params // (world * 64)) * world * 64)
        buf = torch.zeros(n, device="cuda")
        for _ in range(3): check(lib.b200xf_allreduce_bucket(model._ctx, ptr(buf), n))
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): check(lib.b200xf_allreduce_bucket(model._ctx, ptr(buf), n))
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        us = float(t.item()) * 1e3
        sweep.append({"mbytes": n * 4 / 1e6, "us": us, "algbw_gbs": n * 4 / (us * 1e-6) / 1e9, "busbw_gbs": n * 4 / (us * 1e-6) / 1e9 * 2 * (world - 1) / world})
        del buf
    out["nccl_allreduce_sweep"] = sweep
    if getattr(model, "_symmetric", None) is not None:
        nv = []
        for mb in (32, 64):
            n = (mb << 18) // (world * 64) * world * 64
            for _ in range(3): check(lib.b200xf_debug_nvls_update(model._ctx, 0, n, 64, 0))
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10): check(lib.b200xf_debug_nvls_update(model._ctx, 0, n, 64, 0))
            e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
            us = float(t.item()) * 1e3
            # per link direction the fused kernel moves the bucket once (every replica's gradients out, everybody's new weights in) plus one shard
            nv.append({"mbytes": n * 4 / 1e6, "us": us, "algbw_gbs": n * 4 / (us * 1e-6) / 1e9, "link_gbs_per_direction": n * 4 * (1 + 1.0 / world) / (us * 1e-6) / 1e9,
                       "frac_of_900": n * 4 * (1 + 1.0 / world) / (us * 1e-6) / 1e9 / 900.0})
        out["nvls_fused_update_sweep"] = nv
    out["busbw_gbs_at_32MB"] = next(x["busbw_gbs"] for x in sweep if abs(x["mbytes"] - 33.554432) < 1)
    out["busbw_frac_of_900"] = out["busbw_gbs_at_32MB"] / 900.0
    out["busbw_if_spread_over_step_gbs"] = out["bus_bytes_per_step"] / (ms_per_step * 1e-3) / 1e9
    if not solo_fits:   # the long configuration fills the device: no room for a second model
        return out
    # the same steps without any collective
    solo = b200xf.Transformer(**hp, max_tokens=B * max(S, T + 1), seed=0)
    solo.trainable_variables; solo.iterations = 1
    for i in range(5): solo.train_step_device(*dev_batches[i % len(dev_batches)], workers_count=world)
    solo.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps): solo.train_step_device(*dev_batches[i % len(dev_batches)], workers_count=world)
    e1.record(); solo.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    solo.close()
    out["ms_per_step_without_collectives"] = float(t.item())
    out["exposed_us"] = (ms_per_step - float(t.item())) * 1e3
    out["busbw_if_spread_over_step_gbs"] = out["bus_bytes_per_step"] / (ms_per_step * 1e-3) / 1e9
    return out


def long_batch_for(cfg, budget_bytes):
    """The long-sequence configuration's batch (BASELINE.json configs[4]: sized to ~150 GB of HBM per GPU) from the library's
    own memory model: 18 B per parameter of state + the activation / scratch arena for B x (T+1) tokens per side."""
    import ctypes as C
    from b200xf import _lib
    lib = _lib.load()

    def need(B):
        hp = _lib.HParams(layers_count=cfg["layers_count"], embedding_size=cfg["embedding_size"], heads_count=cfg["heads_count"],
                          feed_forward_hidden_size=cfg["feed_forward_hidden_size"], input_vocabulary_size=cfg["input_vocabulary_size"],
                          target_vocabulary_size=cfg["target_vocabulary_size"], maximum_input_length=max(1000, cfg["S"]),
                          maximum_target_length=max(1000, cfg["T"] + 1), droput_rate=0.0, adam_beta_1=0.9, adam_beta_2=0.98, adam_epsilon=1e-9,
                          warmup_steps=4000, max_tokens=B * (cfg["T"] + 1))
        a, b = C.c_int64(), C.c_int64()
        _lib.check(lib.b200xf_device_bytes(C.byref(hp), C.byref(a), C.byref(b)))
        return a.value, b.value
    B = 16
    while sum(need(B + 16)) <= budget_bytes:
        B += 16
    st, ws = need(B)
    return B, {"budget_gb": budget_bytes / 1e9, "sentences_per_gpu": B, "state_gb": st / 1e9, "workspace_gb": ws / 1e9,
               "gb_per_sentence": (need(B)[1] - need(B - 16)[1]) / 16 / 1e9,
               "model": "b200xf_device_bytes: 18 B/parameter (fp32 weights, gradients, Adam m, v, bf16 copy) + arena(max_tokens = B x 1025)"}


# ---------------------------------------------------------------------------------------------- main arm
def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    import numpy as np
    import torch
    import b200xf
    world = int(os.environ.get("WORLD_SIZE", 1)); rank = int(os.environ.get("RANK", 0)); local = int(os.environ.get("LOCAL_RANK", 0))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the b200xf path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg = dict(b200xf.CONFIGS[args.config])
    S, T = cfg["S"], cfg["T"]
    B = args.batch or cfg["B"]
    memory_model = None
    if args.config == "long" and not args.batch:
        B, memory_model = long_batch_for(cfg, args.hbm_gb * 1e9)
    cfg["B"] = B
    hp = dict(layers_count=cfg["layers_count"], embedding_size=cfg["embedding_size"], heads_count=cfg["heads_count"],
              feed_forward_hidden_size=cfg["feed_forward_hidden_size"], input_vocabulary_size=cfg["input_vocabulary_size"],
              target_vocabulary_size=cfg["target_vocabulary_size"], maximum_input_length=max(1000, S), maximum_target_length=max(1000, T + 1),
              droput_rate=0.0)
    model = b200xf.Transformer(**hp, max_tokens=B * max(S, T + 1), seed=0)
    strategy = b200xf.B200Strategy(rank, world)
    model.trainable_variables  # build
    strategy.attach(model)
    model.iterations = 1  # lr(0) == 0 would make the first update a no-op
    lib = b200xf._lib.load()

    nbatch = 8
    host = [b200xf.synthetic_batch(B, S, T + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=1234 + 1000 * rank + i, ragged=args.ragged)
            for i in range(nbatch)]
    pinned = [(torch.from_numpy(a).pin_memory(), torch.from_numpy(b).pin_memory()) for a, b in host]
    dev = [(a.cuda(), b.cuda()) for a, b in pinned]
    W = max(args.warmup, 3); Ksteps = args.steps

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident leg
    sampler = ClockSampler(local)
    if rank == 0: sampler.start()
    for i in range(W):
        out = model.train_step_device(*dev[i % nbatch], workers_count=world)
    model.synchronize(); barrier()
    l0 = lib.b200xf_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    wall0 = time.time()
    e0.record()
    for i in range(Ksteps):
        out = model.train_step_device(*dev[(W + i) % nbatch], workers_count=world)
    e1.record()
    model.synchronize(); barrier()
    sampler.window(wall0, time.time())
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = lib.b200xf_launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    loss = float(out[0].item())
    tokens_per_step = B * T * world
    value = tokens_per_step * Ksteps / (ms / 1e3)

    # ---- end-to-end leg: host (pinned) token buffers through b200xf_train_step_host
    for i in range(2):
        model.train_step_host(*pinned[i % nbatch], workers_count=world)
    barrier()
    t0 = time.perf_counter()
    for i in range(Ksteps):
        res = model.train_step_host(*pinned[i % nbatch], workers_count=world)  # returns after the D2H of the scalars
    barrier()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    e2e = {"value": tokens_per_step * Ksteps / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int((B * S + B * (T + 1)) * 8),
           "d2h_bytes_per_step": 24, "ms_per_step": e2e_ms / Ksteps, "timer": "host wall clock around K synchronous calls, max over ranks"}

    n_params = int(lib.b200xf_param_total(model._ctx))
    # ---- parity of the measured configuration (every rank; outside the timed regions)
    parity = None
    if not args.no_parity:
        parity = parity_block(model, args.config, cfg, hp, host[0], world, rank, dist if world > 1 else None)
    allreduce = None
    if world > 1 and not args.no_allreduce_probe:
        allreduce = allreduce_block(model, hp, B, S, T, world, rank, dist, ms / Ksteps, n_params, dev, min(Ksteps, 30), solo_fits=args.config != "long")
    if rank != 0:
        model.close()
        if world > 1: dist.destroy_process_group()
        return
    peak_burst, peak_sust, peak_hbm, peak_src = measured_peaks()
    flops_step = b200xf.step_flops(hp["layers_count"], hp["embedding_size"], hp["feed_forward_hidden_size"], hp["target_vocabulary_size"], B * S, B * T, S, T)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": Ksteps, "warmup": W, "ms_per_step": ms / Ksteps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": config_dict(args.config, cfg, B, world, args.ragged), "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
            "loss": loss, "step_tflops_algorithmic": flops_step / 1e12,
            "step_tflops_achieved_per_gpu": flops_step / (ms / Ksteps / 1e3) / 1e12,
            "step_frac_of_gemm_roofline": (flops_step / (ms / Ksteps / 1e3)) / (peak_sust * 1e12),
            "step_frac_of_nominal_2250tf": (flops_step / (ms / Ksteps / 1e3)) / 2.25e15, "peak_source": peak_src}
    if memory_model is not None:
        line["memory_model"] = memory_model
        line["device_memory_allocated_gb"] = torch.cuda.mem_get_info()[1] / 1e9 - torch.cuda.mem_get_info()[0] / 1e9
    if parity is not None:
        line["parity"] = parity
    if allreduce is not None:
        line["allreduce"] = allreduce
    model.close()
    if not args.no_roofline:
        rB = min(B, 64) if args.config == "long" else B   # long: the isolated replays use 64 sentences (65 536 tokens) per launch set
        gflops, gms, nl = gemm_roofline(cfg, rB * S, rB * T)
        ach = gflops / (gms / 1e3) / 1e12
        line["roofline"] = {"bound": "tensor", "kernel": "b200xf::gemm_kernel (tcgen05)", "achieved": ach, "peak": peak_burst, "unit": "TFLOP/s", "frac": ach / peak_burst,
                            "traffic": None, "launches_per_step": nl, "avg_launch_us": gms * 1e3 / nl, "flops_per_launch": gflops / nl,
                            "gemm_ms_per_step": gms, "peak_kind": "burst (the kernel is timed alone, 3 replays of ~2 ms)",
                            "how": "all GEMM launches of one step with the step's own epilogues (bias, ReLU, residual add, ReLU mask, fp32 reduce-add) "
                                   "captured in a CUDA graph and replayed back to back, CUDA events on the launching stream, mean of 3 replays"}
        if args.config in ("tutorial", "base", "big"):
            iflops, ibusy_ms, inl, span_ms = gemm_in_step(hp, B, S, T, B * max(S, T + 1))
            iach = iflops / (ibusy_ms / 1e3) / 1e12
            line["roofline"]["in_step"] = {"achieved": iach, "peak": peak_sust, "frac": iach / peak_sust, "launches": inl, "sum_launch_durations_ms": ibusy_ms,
                                           "avg_launch_us": ibusy_ms * 1e3 / max(inl, 1), "first_to_last_gemm_ms": span_ms,
                                           "gemm_share_of_step": ibusy_ms / (ms / Ksteps), "peak_kind": "sustained (kernels timed inside a long step)",
                                           "how": "%globaltimer stamps written by every GEMM CTA of one graph-replayed train step; a launch lasts from its median "
                                                  "CTA's dependency-resolved stamp to its last CTA's exit; weight-gradient GEMMs run concurrently on a side stream, so the sum can exceed their share of the wall clock"}
        if rB != B:
            line["roofline"]["note"] = f"isolated replays at {rB} sentences ({rB * T} tokens) per launch set; the timed step runs {B} sentences"
        line["roofline"]["secondary"] = secondary_rooflines(cfg, rB * T, n_params, peak_hbm)
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_block(args.config, cfg, quick=args.steps < 50)
    print(json.dumps(line), flush=True)
    if world > 1: dist.destroy_process_group()


if __name__ == "__main__":
    main()
