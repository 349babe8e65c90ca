"""bench.py --impl reference on CPU: the JSON line keeps the driver's contract, describes the run it
actually made, and the process never imports the product package or maps libpynama_b200.so
(VERDICT round 1, weak #4).  Run at a small mesh so that the CPU suite stays short; the arm's code path
(C port of the oracle built by oracle/Makefile, all host threads) is the one the full-size run takes."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WRAP = r"""
import atexit, runpy, sys
def report():
    maps = open('/proc/self/maps').read()
    sys.stderr.write('PRODUCT_IMPORTED=%d PRODUCT_SO_MAPPED=%d TORCH_IMPORTED=%d\n' % (
        any(m == 'pynama_b200' or m.startswith('pynama_b200.') for m in sys.modules),
        'libpynama_b200' in maps, 'torch' in sys.modules))
atexit.register(report)
sys.argv = ['bench.py'] + sys.argv[1:]
runpy.run_path('bench.py', run_name='__main__')
"""


def _run(args, env_extra=None, timeout=600):
    env = dict(os.environ)
    env.update(env_extra or {})
    r = subprocess.run([sys.executable, "-c", WRAP] + args, cwd=ROOT, env=env, capture_output=True, text=True,
                       timeout=timeout)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.strip().split("\n") if ln.startswith("{")]
    assert len(lines) == 1, r.stdout
    return json.loads(lines[0]), r.stderr


@pytest.fixture(scope="module")
def line():
    if not os.path.exists(os.path.join(ROOT, "oracle", "libckle.so")) and \
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], capture_output=True).returncode != 0:
        pytest.skip("C port of the oracle does not build here")
    return _run(["--impl", "reference", "--nelem", "4", "--ngl", "3", "--steps", "2", "--warmup", "1"],
                {"PYN_CPU_BUDGET_S": "30"})


def test_reference_line_keeps_the_contract(line):
    rec, _ = line
    assert rec["impl"] == "reference" and rec["metric"] == "kle_timestep_throughput" and rec["unit"] == "DOF/s"
    assert rec["higher_is_better"] is True and rec["vs_baseline"] is None and rec["dtype"] == "f64"
    assert rec["n_gpus"] == 1 and rec["gpu_launches"] == 0 and rec["scaling"] == "strong"
    # the line reports the steps it ran (here: all that were asked for) next to the request
    assert rec["steps"] == rec["steps_requested"] == 2 and rec["warmup"] == rec["warmup_requested"] == 1
    assert abs(rec["value"] - rec["config"]["dofs"] / rec["ms_per_step"] * 1e3) <= 1e-6 * rec["value"]
    wl = rec["config"]["workload"]
    assert rec["config"]["dofs"] == 9 ** 3 * 3 and "4^3 elements, ngl=3" in wl and "NOT the configs[1] size" in wl
    cb = rec["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] == rec["value"] and cb["cores"] >= 1 and cb["sample"]
    e = rec["e2e"]
    assert e["value"] == rec["value"] and e["unit"] == rec["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    ck = rec["checksum"]
    assert ck["vort_l2"] > 0 and ck["vel_l2"] > 0 and ck["after_steps"] >= 1


def test_reference_arm_never_loads_the_product(line):
    _, err = line
    assert "PRODUCT_IMPORTED=0 PRODUCT_SO_MAPPED=0" in err, err[-500:]


def test_reference_arm_cuts_the_run_to_its_budget_and_says_so():
    """with a budget too small for the request the line reports the steps really timed, not the request"""
    rec, _ = _run(["--impl", "reference", "--nelem", "6", "--ngl", "3", "--steps", "400", "--warmup", "3"],
                  {"PYN_CPU_BUDGET_S": "2"})
    assert rec["steps_requested"] == 400 and 1 <= rec["steps"] < 400
    assert rec["warmup"] <= 3
    assert abs(rec["value"] - rec["config"]["dofs"] / rec["ms_per_step"] * 1e3) <= 1e-6 * rec["value"]
