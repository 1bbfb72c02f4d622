"""bench.py's reference arm (the CPU leg of the contract) runs here, without a GPU: one JSON line with the
keys the driver reads, a bounded sample, and - under torchrun - work on rank 0 only."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_bench(*args, env=None):
    e = dict(os.environ)
    e.pop("RANK", None); e.pop("WORLD_SIZE", None); e.pop("LOCAL_RANK", None)
    e.update(env or {})
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, env=e,
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return [ln for ln in out.stdout.splitlines() if ln.startswith("{")]


def test_reference_arm_json_contract():
    lines = run_bench("--impl", "reference", "--steps", "2", "--warmup", "1")
    assert len(lines) == 1
    r = json.loads(lines[0])
    assert r["impl"] == "reference" and r["metric"] == "llama2_7b_pq_decode_tokens_per_s" and r["unit"] == "tokens/s"
    assert r["steps"] == 2 and r["warmup"] == 1 and r["n_gpus"] == 1 and r["higher_is_better"] is True
    assert r["value"] > 0 and r["ms_per_step"] > 0 and r["vs_baseline"] is None and r["data"] == "synthetic"
    assert r["config"]["workload"].startswith("llama-2-7b") and "model" not in r["config"]
    cb = r["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == r["value"] and "aqlm.py" in cb["sample"]
    assert r["e2e"] == {"value": r["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_without_work():
    lines = run_bench("--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                      env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert lines == []


def test_our_arm_has_no_cpu_fallback():
    e = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        e.pop(k, None)
    e["CUDA_VISIBLE_DEVICES"] = ""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1", "--no-cpu",
                          "--no-extras"], cwd=ROOT, env=e, capture_output=True, text=True, timeout=600)
    assert out.returncode != 0 and "no CPU fallback" in out.stderr
    assert not [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
