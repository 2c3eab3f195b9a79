"""The C++ drivers (apps/*.cpp over proton_host.hpp + the C ABI): they build, refuse to run without a GPU (no CPU
fallback), and on a B200 print the reference's golden numbers."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "apps", "bin")


def _build():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "apps"), "-s"])


def _have_gpu():
    import torch
    return torch.cuda.is_available()


def test_drivers_build_and_fail_loudly_without_gpu():
    _build()
    for exe in ("convergence_test", "cuthho_square", "obstacle"):
        assert os.path.exists(os.path.join(BIN, exe))
    if _have_gpu():
        pytest.skip("GPU present")
    r = subprocess.run([os.path.join(BIN, "obstacle"), "-k", "1", "-N", "8"], capture_output=True, text=True)
    assert r.returncode == 2 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_obstacle_driver_prints_reference_error(golden):
    _build()
    r = subprocess.run([os.path.join(BIN, "obstacle"), "-k", "1", "-N", "16"], capture_output=True, text=True, check=True)
    err = float(re.search(r"Error: ([0-9.eE+-]+)", r.stdout).group(1))
    ref = golden["obstacle"]["16"]["k1"]
    assert abs(err - ref) / ref < 1e-5


@pytest.mark.gpu
def test_cuthho_square_driver_prints_reference_error(golden, tmp_path):
    _build()
    r = subprocess.run([os.path.join(BIN, "cuthho_square"), "-f", "-k", "1", "-M", "10", "-N", "10", "-r", "4"], capture_output=True, text=True, check=True, cwd=tmp_path)
    err = float(re.search(r"H1-error : ([0-9.eE+-]+)", r.stdout).group(1))
    ref = golden["cuthho_fictdom"]["k1_N10_r4"]
    assert abs(err - ref) / ref < 5e-6


@pytest.mark.gpu
def test_cuthho_square_interface_driver_prints_reference_error(golden, tmp_path):
    _build()
    r = subprocess.run([os.path.join(BIN, "cuthho_square"), "-i", "-k", "1", "-M", "20", "-N", "20", "-r", "4"], capture_output=True, text=True, check=True, cwd=tmp_path)
    err = float(re.search(r"Energy-norm absolute error:\s+([0-9.eE+-]+)", r.stdout).group(1))
    ref = golden["cuthho_interface"]["k1_N20_r4"]
    assert abs(err - ref) / ref < 1e-5
    assert (tmp_path / "cuthho_cg_hist.dat").exists()
    rows = (tmp_path / "interface_uT.dat").read_text().splitlines()
    assert len(rows) > 20 * 20 * 4 * 6 and len(rows[0].split()) == 3


@pytest.mark.gpu
def test_convergence_test_driver_rates(tmp_path):
    _build()
    cfg = tmp_path / "cfg.lua"
    cfg.write_text("-- flat config, the reference's keys\ndeg_min = 1\ndeg_max = 2\nmin_N = 8\nsteps = 3\nprecond = true\nstab_hho = true\n")
    r = subprocess.run([os.path.join(BIN, "convergence_test"), str(cfg)], capture_output=True, text=True, check=True, cwd=tmp_path)
    rates = re.findall(r"L2 [0-9.e+-]+ \(([0-9.]+)\)\s+energy [0-9.e+-]+ \(([0-9.]+)\)", r.stdout)
    assert len(rates) == 4
    # HHO(k+1,k): L2 order k+2, energy order k+1 (convergence_test.cpp:312-325)
    for (l2, en), k in zip(rates, (1, 1, 2, 2)):
        assert abs(float(l2) - (k + 2)) < 0.25 and abs(float(en) - (k + 1)) < 0.15
    # the reference's output files (convergence_test.cpp:155-161, 232-242, 250-252, 311-312)
    its = [int(v) for v in re.findall(r"CG its (\d+)", r.stdout)]
    for k in (1, 2):
        rows = [l.split() for l in (tmp_path / f"hho_history_precond_{k}.txt").read_text().splitlines()]
        assert len(rows) == 3 and all(len(x) == 3 for x in rows) and abs(float(rows[0][0]) - 2 ** 0.5 / 8) < 1e-6
        for i, N in enumerate((8, 16, 32)):
            hist = [float(v) for v in (tmp_path / f"cg_history_precond_{N}_{k}.txt").read_text().split()]
            assert hist[0] == 1.0 and hist[-1] < 1e-12 and len(hist) == its[(k - 1) * 3 + i] + 2
            cd = k + 1; nq = ((2 * cd + 2) // 2) ** 2           # tensor Gauss points of degree 2*cd per cell
            dat = (tmp_path / f"solution_N_{N}_k_{k}.dat").read_text().splitlines()
            assert len(dat) == N * N * nq and len(dat[0].split()) == 3


@pytest.mark.gpu
def test_reference_loops_uncut(tmp_path):
    """the reference's own per-cell assembly loop (convergence_test.cpp:199-247) compiled against proton_reference_api.hpp"""
    _build()
    for mode, k in (("uncut", 1), ("uncut-naive", 2)):
        r = subprocess.run([os.path.join(BIN, "reference_loops"), mode, "16", str(k)], capture_output=True, text=True, check=True, cwd=tmp_path)
        m = re.search(r"unknowns (\d+) L2 ([0-9.e+-]+) energy ([0-9.e+-]+) take_local_data_vs_sol ([0-9.e+-]+)", r.stdout)
        assert m, r.stdout + r.stderr
        assert int(m.group(1)) == (2 * 16 * 15) * (k + 1)
        assert float(m.group(4)) < 1e-13
        if mode == "uncut":      # direct-solve errors of the restatement at N=16, k=1
            assert abs(float(m.group(2)) - 3.5100744911e-4) < 1e-8 * 3.5e-4 * 100 and abs(float(m.group(3)) - 8.6421724735e-3) < 1e-6 * 8.6e-3


@pytest.mark.gpu
def test_reference_loops_cut(golden, tmp_path):
    """run_cuthho_fictdom's loop (cuthho_square.cpp:870-900) through the per-cell API: the reference's table value"""
    _build()
    r = subprocess.run([os.path.join(BIN, "reference_loops"), "cut", "10", "1", "4"], capture_output=True, text=True, check=True, cwd=tmp_path)
    m = re.search(r"cut cells (\d+) H1-error ([0-9.e+-]+)", r.stdout)
    assert m, r.stdout + r.stderr
    assert int(m.group(1)) == 28
    ref = golden["cuthho_fictdom"]["k1_N10_r4"]
    assert abs(float(m.group(2)) - ref) / ref < 5e-6
