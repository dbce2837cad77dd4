#!/usr/bin/env python
"""Benchmark of the data-parallel Transformer train step (BASELINE.json metric: target tokens/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config base|tutorial|big|long] [--impl reference]

One process per GPU (torchrun sets RANK/LOCAL_RANK/WORLD_SIZE for N > 1).  A "step" is one train_step
(forward + masked loss + backward + gradient all-reduce + Adam) over one synthetic batch.  Prints ONE JSON line.
  value    device-timed whole-job target tokens/s, token batches already resident in HBM
  e2e      the same metric through the HOST-buffer C-ABI call (H2D of the tokens and D2H of the loss inside the timed region)
  roofline the tcgen05 GEMM kernel: algorithmic FLOPs of the step's GEMM launches / their CUDA-event time, vs the measured bf16 peak
  cpu_baseline  the CPU oracle (a torch fp32 restatement of the reference; TensorFlow itself is not installable here) on the host cores
`--impl reference` times that CPU restatement alone, same metric/config, bounded sample per step.
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
def cpu_reference(name, cfg, steps, warmup, sample_sentences):
    """The reference's train_step restated on the CPU (oracle/), all host threads, on a bounded sample of the workload."""
    import torch
    from oracle import transformer_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hp = dict(layers_count=cfg["layers_count"], embedding_size=cfg["embedding_size"], heads_count=cfg["heads_count"],
              feed_forward_hidden_size=cfg["feed_forward_hidden_size"], input_vocabulary_size=cfg["input_vocabulary_size"],
              target_vocabulary_size=cfg["target_vocabulary_size"], maximum_input_length=max(1000, cfg["S"]),
              maximum_target_length=max(1000, cfg["T"] + 1), droput_rate=0.0)
    model = O.OracleTransformer(**hp).load(O.init_parameters(hp["layers_count"], hp["embedding_size"], hp["feed_forward_hidden_size"],
                                                             hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=0))
    model.iterations = 1
    Bs = sample_sentences
    times = []
    for i in range(warmup + steps):
        inp, tgt = O.synthetic_batch(Bs, cfg["S"], cfg["T"] + 1, hp["input_vocabulary_size"], hp["target_vocabulary_size"], seed=1234 + i, ragged=False)
        t0 = time.perf_counter()
        O.train_step(model, inp, tgt, 1)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    per_step = sum(times) / len(times)
    return dict(value=Bs * cfg["T"] / per_step, unit=UNIT, cores=cores, kind="port",
                sample=f"{Bs} of the workload's sentences per step ({Bs * cfg['T']} target tokens), {len(times)} timed steps, "
                       f"torch-CPU fp32 restatement of the reference (TensorFlow not installable here)"), per_step


def run_reference(args):
    import b200xf  # only for the config table (pure Python import; no GPU work)
    cfg = dict(b200xf.CONFIGS[args.config])
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    sample = {"tutorial": 64, "base": 4, "big": 2, "long": 1}[args.config]
    steps = max(1, min(args.steps, 5)); warmup = max(1, min(args.warmup, 1))
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


# ---------------------------------------------------------------------------------------------- GEMM roofline leg
def gemm_shapes(cfg, Ns, Nt):
    """Every GEMM launch of one train step as (M, N, K, layout): forward y=xW, dgrad dx=dyW^T, wgrad dW=x^T dy."""
    L, d, dff, V = cfg["layers_count"], cfg["embedding_size"], cfg["feed_forward_hidden_size"], cfg["target_vocabulary_size"]
    dense = []  # (tokens, K_in, N_out)
    for _ in range(L):
        dense += [(Ns, d, 3 * d), (Ns, d, d), (Ns, d, dff), (Ns, dff, d)]
    dense += [(Ns, d, L * 2 * d)]
    for _ in range(L):
        dense += [(Nt, d, 3 * d), (Nt, d, d), (Nt, d, d), (Nt, d, d), (Nt, d, dff), (Nt, dff, d)]
    dense += [(Nt, d, V)]
    shapes = []
    for tok, K, N in dense:
        shapes += [(tok, N, K, "fwd"), (tok, K, N, "dgrad"), (K, N, tok, "wgrad")]
    return shapes


def gemm_roofline(cfg, Ns, Nt, reps=3):
    import torch
    from b200xf import kernels as Kn
    shapes = gemm_shapes(cfg, Ns, Nt)
    bufs = {}
    LAY = {"fwd": (False, True), "dgrad": (False, False), "wgrad": (True, True)}

    def operands(M, N, K, lay):
        key = (M, N, K, lay)
        if key not in bufs:
            a_mn, b_mn = LAY[lay]
            pad = lambda n: (n + 7) // 8 * 8
            a = torch.randn((K, pad(M)) if a_mn else (M, pad(K)), device="cuda").to(torch.bfloat16)
            b = torch.randn((K, pad(N)) if b_mn else (N, pad(K)), device="cuda").to(torch.bfloat16)
            a = a[:, :M] if a_mn else a[:, :K]; b = b[:, :N] if b_mn else b[:, :K]
            out = torch.zeros((M, pad(N)), device="cuda", dtype=torch.float32 if lay == "wgrad" else torch.bfloat16)[:, :N]
            bufs[key] = (a, b, out, a_mn, b_mn)
        return bufs[key]

    def run_all():
        for M, N, K, lay in shapes:
            a, b, out, a_mn, b_mn = operands(M, N, K, lay)
            Kn.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, accumulate=(lay == "wgrad"))
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
    flops = sum(2.0 * M * N * K for M, N, K, _ in shapes)
    return flops, ms, len(shapes)


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        j = json.loads(p.read_text())
        return float(j.get("bf16_tflops_sustained", j.get("bf16_tflops"))), float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json, sustained)"
    return 1400.0, 6650.0, "fallback (B200_PROFILING.md)"


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
    B = args.batch or cfg["B"]; S, T = cfg["S"], cfg["T"]
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

    if rank != 0:
        model.close()
        if world > 1: dist.destroy_process_group()
        return
    peak_tf, peak_hbm, peak_src = measured_peaks()
    flops_step = b200xf.step_flops(hp["layers_count"], hp["embedding_size"], hp["feed_forward_hidden_size"], hp["target_vocabulary_size"], B * S, B * T, S, T)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": Ksteps, "warmup": W, "ms_per_step": ms / Ksteps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": config_dict(args.config, cfg, B, world, args.ragged), "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
            "loss": loss, "step_tflops_algorithmic": flops_step / 1e12,
            "step_frac_of_gemm_roofline": (flops_step * world / (ms / Ksteps / 1e3)) / (peak_tf * 1e12 * world), "peak_source": peak_src}
    if not args.no_roofline:
        gflops, gms, nl = gemm_roofline(cfg, B * S, B * T)
        ach = gflops / (gms / 1e3) / 1e12
        line["roofline"] = {"bound": "tensor", "kernel": "b200xf::gemm_kernel (tcgen05)", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach / peak_tf,
                            "traffic": None, "launches_per_step": nl, "avg_launch_us": gms * 1e3 / nl, "flops_per_launch": gflops / nl,
                            "gemm_ms_per_step": gms, "gemm_share_of_step": gms / (ms / Ksteps),
                            "how": "all GEMM launches of one step captured in a CUDA graph and replayed back to back, CUDA events on the launching stream, mean of 3 replays"}
    if world == 1 and not args.no_cpu_baseline:
        sample = {"tutorial": 64, "base": 4, "big": 2, "long": 1}[args.config]
        cb, _ = cpu_reference(args.config, cfg, 2, 1, sample)
        line["cpu_baseline"] = cb
    print(json.dumps(line), flush=True)
    model.close()
    if world > 1: dist.destroy_process_group()


if __name__ == "__main__":
    main()
