"""bench.py's reference arm (the reference's own MultigridSmoother.step from the staged copy oracle/_ref, or the NumPy
port where that copy is absent) runs without a GPU: its JSON line carries the keys the driver reads, and the single-cavity
step timer used for the CPU side of time-to-tolerance returns a sane figure."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                          "--cpu-cores", "2", "--N", "16"], stdout=subprocess.PIPE, env=env, text=True, check=True, timeout=120).stdout
    lines = [ln for ln in out.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["unit"] == "cavity-pseudo-steps/s" and d["dtype"] == "f64" and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["gpu_launches"] == 0
    staged = os.path.isdir(os.path.join(ROOT, "oracle", "_ref", "src", "solvers")) or os.path.isdir("/root/reference/src/solvers")
    assert d["cpu_baseline"]["kind"] == ("reference" if staged else "port")
    assert d["cpu_baseline"]["cores"] == 2 and d["cpu_baseline"]["value"] == d["value"]
    if not staged:
        assert "not used" in d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_do_nothing():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], stdout=subprocess.PIPE, env=env, text=True, timeout=60)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_single_cavity_step_timer():
    sys.path.insert(0, ROOT)
    import bench

    t, kind = bench.cpu_step_seconds(12, seconds=0.3)
    assert 1e-6 < t < 0.1 and kind in ("reference", "port")
    t, kind = bench.cpu_step_seconds(12, seconds=0.3, port=True)
    assert 1e-6 < t < 0.1 and kind == "port"


def test_port_flag_times_the_oracle_port():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--cpu-cores", "1", "--N", "12", "--cpu-port"], stdout=subprocess.PIPE, env=env, text=True, check=True,
                         timeout=120).stdout
    d = json.loads(out.strip().splitlines()[-1])
    assert d["cpu_baseline"]["kind"] == "port" and "--cpu-port" in d["cpu_baseline"]["sample"]
