"""CPU: the part of bench.py's contract that runs without a GPU - the reference arm
(`--impl reference`: the cv2 restatement of the reference on the host cores). One JSON line on
stdout with the keys the driver reads; under a multi-rank launch only rank 0 works and prints."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None):
    env = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    env.update(extra_env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                           "--steps", "1", "--warmup", "0", "--ref-frames", "4"],
                          capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)


def test_reference_arm_prints_one_line_with_the_contract_keys():
    r = _run()
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    assert d["metric"].startswith("stereo frames/s") and d["unit"] == "frames/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["vs_baseline"] is None and d["gpu_launches"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    # "reference": oracle/_ref (the reference's own compiled front end) was usable; "port": it was not
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_frontend.so")) or os.path.exists("/root/reference/src"):
        assert cb["kind"] == "reference", cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    r = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip() == ""


def test_reference_arm_falls_back_to_the_port_without_the_compiled_reference():
    """a machine with neither the reference tree nor the built library (played through the two
    environment overrides of oracle/build_ref.py): same line, kind "port", the reason stated"""
    r = _run({"ORACLE_REF_SRC": "/nonexistent", "ORACLE_REF_OUT": "/nonexistent"})
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([l for l in r.stdout.splitlines() if l.strip()][-1])
    assert d["impl"] == "reference" and d["cpu_baseline"]["kind"] == "port" and d["value"] > 0
    assert "not usable here" in d["cpu_baseline"]["sample"]
