"""bench.py contract on CPU: the reference arm (the reference's own modules from oracle/_ref, or the
oracle port when they are absent, timed on the host cores) prints the agreed JSON line, and our arm has no CPU fallback (it must fail, not quietly time something else)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, env=None):
    e = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, env=e,
                          capture_output=True, text=True, timeout=600)


def test_reference_arm_prints_the_contract_line():
    r = _run("--impl", "reference", "--config", "tiny-256", "--steps", "1", "--warmup", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "images/s" and line["higher_is_better"] is True
    assert line["metric"].startswith("concept-heatmap images/sec") and line["value"] > 0
    assert line["n_gpus"] == 1 and line["steps"] == 1 and line["warmup"] == 1 and line["vs_baseline"] is None
    cb = line["cpu_baseline"]
    from oracle import reference_modules as rm
    assert cb["kind"] == ("reference" if rm.available() else "port")
    assert cb["cores"] >= 1 and cb["sample"] and cb["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "images/s", "h2d_bytes_per_step": 0,
                           "d2h_bytes_per_step": 0}
    assert "tiny-256" in line["config"]["workload"] and "model" not in line["config"]


def test_reference_arm_other_ranks_do_no_work():
    r = _run("--impl", "reference", "--config", "tiny-256", "--steps", "1", "--warmup", "1", "--gpus", "2",
             env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip() == ""


def test_product_arm_has_no_cpu_fallback():
    r = _run("--config", "tiny-256", "--steps", "1", "--warmup", "1")
    assert r.returncode != 0
    assert not any(l.startswith("{") for l in r.stdout.splitlines())      # no bench line without a GPU
