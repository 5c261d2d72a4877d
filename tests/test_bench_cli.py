"""bench.py on a machine without a GPU: the command line the driver uses parses, the
workloads of BASELINE.json exist with the parameter lists it quotes, and the reference
arm (the one leg of bench.py that needs no GPU) prints the contract's JSON line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _parse(argv):
    import bench
    old = sys.argv
    sys.argv = ["bench.py"] + argv
    try:
        return bench.parse_args()
    finally:
        sys.argv = old


def test_driver_command_lines_parse():
    a = _parse([])
    assert (a.gpus, a.impl, a.workload) == (1, "b200", os.environ.get("RAILS_BENCH_WORKLOAD", "C3"))
    assert a.warmup >= 3                                  # timing rule: W >= 3 warm-up steps
    a = _parse(["--gpus", "8", "--steps", "20", "--warmup", "5"])
    assert (a.gpus, a.steps, a.warmup) == (8, 20, 5)
    a = _parse(["--impl", "reference", "--gpus", "2", "--steps", "4", "--warmup", "3"])
    assert a.impl == "reference"
    a = _parse(["--workload", "C5", "--gpus", "8", "--windows", "3"])
    assert a.workload == "C5" and not a.no_continuation
    a = _parse(["--parity-full", "--workload", "C4", "--parity-cpu-grid", "40"])
    assert a.parity_full and a.parity_cpu_grid == 40


def test_workloads_are_the_baseline_configs():
    from rails_b200 import problems
    C = problems.CONFIGS
    assert C["C1"][:4] == ("laplacian", 64, 2, 1)
    assert C["C2"][:4] == ("laplacian", 1024, 2, 4) and C["C2"][4]["Restart size"] == 200
    assert C["C3"][:4] == ("laplacian", 256, 3, 1) and C["C3"][4]["Use mass matrix"] == 1
    # config 4 with BASELINE's own list: no size cap (the capped variant is C4CAP)
    assert C["C4"][:4] == ("convdiff", 256, 3, 16)
    assert "Restart size" not in C["C4"][4] and C["C4"][4]["Expand size"] == 16
    assert C["C4CAP"][4]["Restart size"] == 192
    assert C["C5"][:4] == ("laplacian", 512, 3, 1)
    for name, (_, _, _, _, P) in C.items():
        assert P["Tolerance"] == 1e-4, name


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--grid", "12",
                          "--steps", "3", "--warmup", "3"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1                                # ONE JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "RAILS outer iterations/s" and d["unit"] == "iterations/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 3 and d["warmup"] == 3
    assert d["value"] > 0 and d["dtype"] == "f64" and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"]
