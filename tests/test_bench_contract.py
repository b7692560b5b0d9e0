"""bench.py contract, the parts that run without a GPU: the reference arm (the CPU oracle port timed on the
host cores) prints ONE JSON line with the keys the driver reads, and the B200 arm refuses to run on the CPU."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def _run(*args, timeout=600):
    return subprocess.run([sys.executable, str(ROOT / "bench.py"), *args], capture_output=True, text=True,
                          timeout=timeout, cwd=str(ROOT))


def test_reference_arm_prints_one_json_line():
    p = _run("--impl", "reference", "--steps", "1", "--warmup", "1", "--cpu-threads", "1", "--history", "256")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "decode tokens/s (whole box)" and d["unit"] == "tokens/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["config"]["workload"].startswith("llama-3-8b")
    # the reference's own host code around the path, timed where its tree is available (here: /root/reference)
    c1 = d["reference_server"]["config1_thread_replay"]
    assert c1["threads"] == 64 and c1["turns"] == 8 and c1["repo"]["messages_in_last_thread"] == 16
    if "unavailable" not in c1["reference"]:
        assert c1["reference"]["messages_in_last_thread"] == 16 and c1["reference"]["turns_per_s"] > 0
        inside = d["reference_server"]["cpu_provider_in_reference_agent"]
        assert inside["tokens"] == inside["threads"] * inside["max_tokens"] and inside["tokens_per_s"] > 0


def test_b200_arm_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        return
    p = _run("--steps", "1", "--warmup", "1", "--no-cpu-baseline", timeout=300)
    assert p.returncode != 0 and "no CPU fallback" in (p.stderr + p.stdout)
