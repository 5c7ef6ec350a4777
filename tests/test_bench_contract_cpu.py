"""bench.py's output contract (one JSON line with the keys the round driver reads), checked on the CPU with the tiny
plumbing model; the reference arm must answer with a JSON line and exit code 0 whether or not it can run here."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*flags, port):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), CUDA_VISIBLE_DEVICES="",
               DSB_BENCH_TINY_SEQ="256")
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *flags], capture_output=True, text=True,
                         env=env, cwd=ROOT, timeout=900)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert lines, out.stdout[-2000:]
    return json.loads(lines[-1])


def test_bench_json_line_has_the_contract_keys():
    d = _run("--tiny", "--steps", "1", "--warmup", "1", port=29631)
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "impl"):
        assert key in d, key
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1 and d["higher_is_better"] is True
    assert d["scaling"] == "weak" and d["unit"] == "tokens/s" and d["impl"] == "ours"
    for key in ("model", "global_batch", "seq_len", "parallelism"):
        assert key in d["config"], key
    for key in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert key in d["e2e"], key
    # tokens and labels of every micro-batch cross the host -> device boundary each step; the loss comes back
    assert d["e2e"]["h2d_bytes_per_step"] == d["config"]["global_batch"] * d["config"]["seq_len"] * 8 * 2
    assert d["e2e"]["d2h_bytes_per_step"] == 4
    assert abs(d["value"] - d["config"]["global_batch"] * d["config"]["seq_len"] / (d["ms_per_step"] / 1e3)) < 1e-6 * d["value"]
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}


def test_reference_arm_always_answers_with_json():
    d = _run("--impl", "reference", "--tiny", "--steps", "1", "--warmup", "1", port=29632)
    assert d["impl"] == "reference" and ("unavailable" in d or "value" in d)
