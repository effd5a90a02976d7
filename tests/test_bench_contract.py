"""The reference arm of bench.py on the host cores (no GPU needed): it must run, print ONE JSON line and carry the keys the
driver reads (`impl`, metric / unit / higher_is_better as in BASELINE.json, `cpu_baseline`, `e2e` with zero copy bytes)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line_on_cpu():
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
           "--rows", "65536", "--cpu-sample-rows", "32768", "--no-full-pass"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout[-2000:]
    j = json.loads(lines[0])
    with open(os.path.join(ROOT, "BASELINE.json")) as f:
        base = json.load(f)
    assert j["impl"] == "reference" and j["higher_is_better"] is True and j["unit"] == "queries/s"
    if isinstance(base.get("metric"), str):
        assert j["metric"].split(",")[0] in base["metric"] or base["metric"].split(",")[0] in j["metric"]
    for key in ("value", "n_gpus", "steps", "warmup", "ms_per_step", "scaling", "dtype", "data", "config"):
        assert key in j, key
    assert j["value"] > 0 and j["ms_per_step"] > 0 and "workload" in j["config"]
    cb = j["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == j["value"] and cb["sample"]
    e = j["e2e"]
    assert e["value"] == j["value"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
