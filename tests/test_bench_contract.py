"""bench.py's reference arm, launched the way the driver launches it (CPU only: it times the C port of the
reference's loop on the host cores): one JSON line from rank 0 with the contract's keys; the other ranks of a
torchrun launch exit 0 without work."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
        "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e")


def _check(line, n_gpus):
    j = json.loads(line)
    for k in KEYS:
        assert k in j, k
    assert j["impl"] == "reference" and j["n_gpus"] == n_gpus and j["value"] > 0 and j["higher_is_better"] is True
    assert j["metric"] == "env_steps_per_sec_incl_perspectives_d7" and j["unit"] == "env-steps/s"
    assert "workload" in j["config"] and j["vs_baseline"] is None
    cb = j["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == j["value"] and cb["sample"]
    assert j["e2e"] == {"value": j["value"], "unit": j["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def _bench_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_reference_arm_single_process(monkeypatch):
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2",
                          "--warmup", "1", "--cpu-envs", "512"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    _check(lines[0], 1)
    # both arms print the SAME config object for the same command line (the driver compares them): it is built by one
    # function from the arguments alone, and the size of the CPU sample is not part of it
    bench = _bench_module()
    monkeypatch.setattr(sys, "argv", ["bench.py", "--steps", "2", "--warmup", "1"])
    a = bench.parse_args()
    assert a.cpu_envs == a.envs == 65536                    # by default the CPU arm walks the batch the GPU arm steps
    assert json.loads(lines[0])["config"] == bench.bench_config(a, 1)
    assert "cpu_envs" in json.loads(lines[0])["cpu_baseline"]


def test_reference_arm_under_torchrun_two_ranks():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
           "--steps", "2", "--warmup", "1", "--cpu-envs", "512"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, res.stdout[-2000:]
    _check(lines[0], 2)
