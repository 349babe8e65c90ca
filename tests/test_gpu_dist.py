"""GPU: the row-partitioned (one process per GPU) path against the oracle.  The single-rank
run of the same script always executes; the 2-rank run needs two visible GPUs."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = os.path.join(ROOT, "tests", "dist_check.py")


def _ngpus():
    out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True)
    return len([l for l in out.stdout.splitlines() if l.startswith("GPU ")]) if out.returncode == 0 else 0


def _run(nproc, port, env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), SCRIPT]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
    assert "OK" in r.stdout


@pytest.mark.parametrize("fischer,ksp", [("0", "cg"), ("4", "cg"), ("0", "gmres"), ("0", "bcgs")])
def test_single_rank_script(fischer, ksp):
    env = dict(os.environ, DIST_FISCHER=fischer, DIST_KSP=ksp)
    r = subprocess.run([sys.executable, SCRIPT], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]


def test_partitioned_run_nccl_and_fischer():
    """2 ranks, NCCL collectives instead of the peer-memory windows, Fischer guess on"""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, 29549, {"PYN_P2P": "0", "DIST_FISCHER": "4"})


@pytest.mark.parametrize("ksp", ["gmres", "bcgs"])
def test_partitioned_run_other_ksp_types(ksp):
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, 29551, {"DIST_KSP": ksp})


@pytest.mark.parametrize("opts", ["ksp_cg_single_reduction=1", "kle_matfree=1",
                                  "ksp_cg_single_reduction=1,kle_matfree=1"])
@pytest.mark.parametrize("p2p", ["1", "0"])
def test_partitioned_run_single_reduction_and_matfree(opts, p2p):
    """2 ranks: single-reduction CG and the matrix-free K, over peer-memory windows and over NCCL"""
    env = {"DIST_OPTS": opts, "PYN_P2P": p2p, "DIST_NGL": "4"}
    if _ngpus() < 2:
        r = subprocess.run([sys.executable, SCRIPT], capture_output=True, text=True, timeout=600,
                           env=dict(os.environ, **env))
        assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
        return
    _run(2, 29553, env)


@pytest.mark.parametrize("nproc", [1, 2, 4])
def test_no_slip_cavity_partitioned(nproc):
    """MatNS (8 matrices), K + Kfs solved with GMRES, tangential wall values: row-partitioned"""
    env = {"DIST_CASE": "cavity", "DIST_NELEM": "3,3,4", "DIST_NGL": "3"}
    if nproc == 1:
        r = subprocess.run([sys.executable, SCRIPT], capture_output=True, text=True, timeout=600,
                           env=dict(os.environ, **env))
        assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
        return
    if _ngpus() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    _run(nproc, 29570 + nproc, env)


@pytest.mark.parametrize("nproc", [1, 2, 4, 8])
def test_cylinder_channel_partitioned(nproc):
    """BASELINE.json configs[3] on a small channel: blocked-cell cylinder, unstructured Gmsh file with the
    two extra wall borders (cylinder-x, cylinder-y), all sides no-slip, RCB partition; fields after two
    RK5BS steps against OracleProblemNS on the global mesh"""
    env = {"DIST_CASE": "cylinder", "DIST_NGL": "3"}
    if nproc == 1:
        r = subprocess.run([sys.executable, SCRIPT], capture_output=True, text=True, timeout=900,
                           env=dict(os.environ, **env))
        assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
        assert "OK" in r.stdout
        return
    if _ngpus() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    _run(nproc, 29580 + nproc, env)


@pytest.mark.parametrize("nproc", [1, 2, 4])
def test_unstructured_mesh_partitioned_by_rcb(nproc):
    """Gmsh input, recursive coordinate bisection over the ranks, same oracle comparison"""
    env = {"DIST_GMSH": "1", "DIST_NELEM": "4,3,4"}
    if nproc == 1:
        r = subprocess.run([sys.executable, SCRIPT], capture_output=True, text=True, timeout=600,
                           env=dict(os.environ, **env))
        assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
        return
    if _ngpus() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    _run(nproc, 29560 + nproc, env)


@pytest.mark.parametrize("nproc", [2, 4, 8])
def test_partitioned_run_matches_oracle(nproc):
    if _ngpus() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    _run(nproc, 29540 + nproc, {"DIST_NELEM": "4,4,4" if nproc < 8 else "4,4,6"})
