"""GPU: drop-in behaviour at the command and link level.

1. The `cg` driver (conjugate-gradient_b200/driver/cg_main.c) keeps the reference's command line
   and output format: its stdout on the config-1 stand-in (2D 5-point 32x32, 30 iterations,
   SURVEY 8d) equals the reference CLI's stdout recorded in tests/golden/config1_cg_stdout.txt
   (%f prints 6 decimals; only the pointer line and the timing line are excluded).
2. The reference's UNMODIFIED cg.c (main, file reader and CG loop), linked against libcgb200.so
   instead of its own mv_ops.c (oracle/_ref/cg_on_b200, built by oracle/Makefile where the
   reference tree is mounted), produces the same output: every operator call of the reference's
   loop then runs as a CUDA kernel.
"""
import json
import os
import subprocess

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

INPUT = os.path.join(H.GOLDEN, "config1_lap2d_32x32.txt")
WANT = os.path.join(H.GOLDEN, "config1_cg_stdout.txt")


def _filtered(stdout):
    return [ln for ln in stdout.splitlines() if not ln.startswith("\tValues:") and not ln.startswith("CG took")]


def _close(got, want, ulps_of_print=1):
    """Line-by-line equality; numeric lines may differ by one unit of the 6th printed decimal
    (the dot products are summed in a different order, north_star tolerance 1e-8 relative)."""
    assert len(got) == len(want)
    for g, w in zip(got, want):
        if g == w:
            continue
        assert abs(float(g) - float(w)) <= 1.0001e-6 * ulps_of_print, (g, w)


def test_cg_driver_matches_reference_cli_output(pkg, gpu_lib):
    out = subprocess.run([pkg.CG_DRIVER, INPUT, "30"], stdout=subprocess.PIPE, text=True, check=True).stdout
    assert out.startswith("CG took approx ")
    _close(_filtered(out), open(WANT).read().splitlines())


def test_cg_driver_suppress_json_and_tolerance(pkg, gpu_lib):
    out = subprocess.run([pkg.CG_DRIVER, INPUT, "5000", "y", "--tol", "1e-10", "--json"], stdout=subprocess.PIPE,
                         text=True, check=True).stdout
    lines = out.splitlines()
    assert lines[0].startswith("CG took approx ") and len(lines) == 2
    j = json.loads(lines[1])
    assert j["status"] == pkg.CONVERGED and j["relres"] <= 1e-10 and 40 < j["iterations"] < 200


def test_cg_driver_fails_loudly_on_bad_input(pkg, gpu_lib, tmp_path):
    r = subprocess.run([pkg.CG_DRIVER, str(tmp_path / "missing.txt"), "3"], stderr=subprocess.PIPE, text=True)
    assert r.returncode != 0 and "Failed to open input file" in r.stderr
    p = tmp_path / "short.txt"
    p.write_text("0,1\n0,1\n1.0\n1.0\n")
    r = subprocess.run([pkg.CG_DRIVER, str(p), "3"], stderr=subprocess.PIPE, text=True)
    assert r.returncode != 0 and "malformed" in r.stderr


def test_cg_driver_accepts_missing_final_newline(pkg, gpu_lib, tmp_path):
    txt = open(INPUT).read().rstrip("\n")          # the reference's reader never terminates on this
    p = tmp_path / "nonl.txt"
    p.write_text(txt)
    out = subprocess.run([pkg.CG_DRIVER, str(p), "30"], stdout=subprocess.PIPE, text=True, check=True).stdout
    _close(_filtered(out), open(WANT).read().splitlines())


def test_cg_driver_device_generated_system(pkg, gpu_lib):
    out = subprocess.run([pkg.CG_DRIVER, "--gen", "lap3d:48", "2000", "y", "--tol", "1e-10", "--json"],
                         stdout=subprocess.PIPE, text=True, check=True).stdout
    j = json.loads(out.splitlines()[-1])
    assert j["status"] == pkg.CONVERGED and j["relres"] <= 1e-10 and j["gpus"] == 1


def test_reference_cg_c_runs_unmodified_on_libcgb200(gpu_lib):
    exe = os.path.join(H.ORACLE_DIR, "_ref", "cg_on_b200")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/cg_on_b200 not built (needs the reference tree at build time)")
    out = subprocess.run([exe, INPUT, "30"], stdout=subprocess.PIPE, text=True, check=True).stdout
    _close(_filtered(out), open(WANT).read().splitlines())


def test_reference_cg_c_unmodified_in_resident_mode(gpu_lib):
    # CGB200_MODE=resident: the reference loop's vectors stay on the device between operator calls
    # (mv_deep_copy copies device to device); only dot_product's return value and the final
    # print_sparse come back to the host.  Same output as the eager run and the CPU build.
    exe = os.path.join(H.ORACLE_DIR, "_ref", "cg_on_b200")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/cg_on_b200 not built (needs the reference tree at build time)")
    env = dict(os.environ, CGB200_MODE="resident")
    out = subprocess.run([exe, INPUT, "30"], stdout=subprocess.PIPE, text=True, check=True, env=env).stdout
    _close(_filtered(out), open(WANT).read().splitlines())


def test_cg_driver_pattern_format_and_variant_flags(pkg, gpu_lib):
    # --pattern: the config-1 stand-in is a 5-point stencil, so the matvec runs in the row-pattern
    # dictionary format; same printed solution.  --variant sr: the single-reduction recurrence.
    out = subprocess.run([pkg.CG_DRIVER, INPUT, "30", "--pattern"], stdout=subprocess.PIPE, text=True, check=True).stdout
    _close(_filtered(out), open(WANT).read().splitlines())
    out = subprocess.run([pkg.CG_DRIVER, "--gen", "lap3d:48", "2000", "y", "--tol", "1e-10", "--json", "--pattern",
                          "--variant", "sr"], stdout=subprocess.PIPE, text=True, check=True).stdout
    j = json.loads(out.splitlines()[-1])
    assert j["status"] == pkg.CONVERGED and j["relres"] <= 1e-10
    r = subprocess.run([pkg.CG_DRIVER, INPUT, "30", "--variant", "nope"], stderr=subprocess.PIPE, text=True)
    assert r.returncode != 0


def test_cg_driver_banded_flag_prints_the_same_solution(pkg, gpu_lib, tmp_path):
    # --banded: a mostly full band is stored as diagonals (8 bytes per entry, no indices); the printed
    # solution is the CSR run's, line for line; a stencil file does not qualify and silently stays on CSR
    s = H.banded(1200, 60)
    p = tmp_path / "band.txt"
    H.write_text_format(s, str(p))
    csr = subprocess.run([pkg.CG_DRIVER, str(p), "12"], stdout=subprocess.PIPE, text=True, check=True).stdout
    dia = subprocess.run([pkg.CG_DRIVER, str(p), "12", "--banded"], stdout=subprocess.PIPE, text=True, check=True).stdout
    _close(_filtered(dia), _filtered(csr))
    out = subprocess.run([pkg.CG_DRIVER, INPUT, "30", "--banded"], stdout=subprocess.PIPE, text=True, check=True).stdout
    _close(_filtered(out), open(WANT).read().splitlines())


def test_cg_driver_dump_writes_a_file_the_readers_accept(pkg, gpu_lib, tmp_path):
    # --gen system -> the reference's text format (and the binary sidecar) -> back through the
    # library's reader: the arrays of the oracle generator; and the file solves like the generated system
    import ctypes as C

    s = H.lap(16, 16)
    txt, binf = tmp_path / "gen.txt", tmp_path / "gen.bin"
    a = subprocess.run([pkg.CG_DRIVER, "--gen", "lap2d:16", "20", "--dump", str(txt)], stdout=subprocess.PIPE, text=True, check=True).stdout
    subprocess.run([pkg.CG_DRIVER, "--gen", "lap2d:16", "20", "y", "--dump-binary", str(binf)], stdout=subprocess.PIPE, check=True)
    L = pkg.lib()
    for path in (txt, binf):
        A, b = L.new_mv_struct(), L.new_mv_struct()
        assert L.cgb200_read_system(str(path).encode(), A, b) == 0, pkg.last_error()
        ac = A.contents
        assert ac.size == s.n and ac.nnz == s.nnz
        assert np.array_equal(np.ctypeslib.as_array(ac.values, shape=(s.nnz,)), s.vals)
        assert np.array_equal(np.ctypeslib.as_array(ac.col_indices, shape=(s.nnz,)), s.cols)
        assert np.array_equal(np.ctypeslib.as_array(ac.row_ptr, shape=(s.n + 1,)), s.row_ptr)
        assert np.array_equal(np.ctypeslib.as_array(b.contents.values, shape=(s.n,)), s.b)
        L.cgb200_free_all(A)
        L.cgb200_free_all(b)
    b_ = subprocess.run([pkg.CG_DRIVER, str(txt), "20"], stdout=subprocess.PIPE, text=True, check=True).stdout
    xa = [l for l in a.splitlines() if l.startswith("\t") and ":" not in l]      # the printed x (mv_ops.c:92-93)
    xb = [l for l in b_.splitlines() if l.startswith("\t") and ":" not in l]
    assert len(xa) == s.n and xa == xb
