"""bench.py's reference arm (the CPU leg of the measurement contract) runs without a GPU: check the JSON line it
prints (same metric / unit / config as our arm, `impl`, `cpu_baseline`, `e2e` with zero copy bytes) and that ranks
other than 0 of a multi-rank launch exit quietly."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra):
    env = dict(os.environ, **env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                          capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)


def test_reference_arm_line():
    out = _run({})
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    sys.path.insert(0, ROOT)
    import bench
    assert d["impl"] == "reference" and d["metric"] == bench.METRIC and d["unit"] == "frames/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["steps"] == 1 and d["n_gpus"] == 1 and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert d["config"]["workload"].startswith("configs[1]")
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and "calcOpticalFlowPyrLK" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    out = _run({"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"})
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_ncu_traffic_record_is_tied_to_source_and_workload(tmp_path, monkeypatch):
    """roofline.traffic comes from the committed ncu capture only while the kernel source (sha1) and the workload are the
    ones it was taken on (no literal constants); anything else reports None with the reason."""
    sys.path.insert(0, ROOT)
    import bench
    import hashlib
    val, src = bench.ncu_traffic("k_lk_track", 64)
    rec = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["k_lk_track"]
    current = hashlib.sha1(open(os.path.join(ROOT, bench.PKG, "csrc", rec["source"]), "rb").read()).hexdigest() == rec["source_sha1"]
    if current:
        assert val == rec["bytes_per_launch"] and val > 2.6e8 and "ncu --set full" in src, src
    else:    # the kernel was edited after the capture: the bench must say so instead of reporting the old figure
        assert val is None and src.startswith("stale")
    assert bench.ncu_traffic("k_lk_track", 32)[0] is None                                   # another batch size: stale
    assert bench.ncu_traffic("k_no_such_kernel", 64)[0] is None
    # a changed kernel source invalidates the record: point ROOT at a copy whose fe_lk.cuh differs by one byte
    fake = tmp_path / "repo"
    (fake / "profiles").mkdir(parents=True)
    (fake / bench.PKG / "csrc").mkdir(parents=True)
    (fake / "profiles" / "ncu_traffic.json").write_text(json.dumps({"k_lk_track": rec}))
    (fake / bench.PKG / "csrc" / "fe_lk.cuh").write_bytes(open(os.path.join(ROOT, bench.PKG, "csrc", "fe_lk.cuh"), "rb").read() + b"\n")
    monkeypatch.setattr(bench, "ROOT", str(fake))
    val2, why = bench.ncu_traffic("k_lk_track", 64)
    assert val2 is None and why.startswith("stale")
