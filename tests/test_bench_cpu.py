"""bench.py's reference arm (the CPU restatement of the reference's train_step; no GPU involved) prints one well-formed JSON line
with the keys the measurement contract names, and its GEMM shape list accounts for the algorithmic FLOPs of a step."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--config", "tutorial", "--steps", "2", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600, cwd=str(ROOT))
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "tokens/s" and line["higher_is_better"] is True
    assert line["metric"].startswith("target tokens/sec") and line["value"] > 0 and line["n_gpus"] == 1
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1 and line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in line["config"] and "model" not in line["config"]


def test_reference_arm_does_nothing_on_other_ranks(monkeypatch):
    import os
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--config", "tutorial", "--gpus", "2", "--steps", "1"],
                         capture_output=True, text=True, timeout=300, cwd=str(ROOT), env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_gemm_shape_list_matches_the_step_flops():
    sys.path.insert(0, str(ROOT))
    import bench
    import b200xf
    for name in ("tutorial", "base", "big"):
        cfg = b200xf.CONFIGS[name]
        Ns, Nt = cfg["B"] * cfg["S"], cfg["B"] * cfg["T"]
        shapes = bench.gemm_shapes(cfg, Ns, Nt)
        L = cfg["layers_count"]
        assert len(shapes) == 3 * (4 * L + 1 + 6 * L + 1)
        gemm_flops = sum(2.0 * M * N * K for M, N, K, _, _ in shapes)
        d = cfg["embedding_size"]
        attention = 3 * L * (4 * Ns * cfg["S"] * d + 4 * Nt * cfg["T"] * d + 4 * Nt * cfg["S"] * d)   # not GEMM launches
        total = b200xf.step_flops(L, d, cfg["feed_forward_hidden_size"], cfg["target_vocabulary_size"], Ns, Nt, cfg["S"], cfg["T"])
        assert abs(gemm_flops + attention - total) / total < 1e-12
