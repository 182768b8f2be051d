"""The C ABI used from plain C99 (tests/c/cabi_smoke.c): include/*.h must compile as C with -pedantic -Werror, the
shared library must link without C++ or Python, and the constructors must fail loudly without a GPU."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "quadruped_pympc_tamols_b200", "lib")


def _build(tmp_path):
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    if not os.path.exists(os.path.join(LIBDIR, "libqpmpc_b200.so")):
        import __graft_entry__
        __graft_entry__.build()
    exe = str(tmp_path / "cabi_smoke")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "c", "cabi_smoke.c"), "-o", exe, "-L", LIBDIR, "-lqpmpc_b200", "-lm",
                    "-Wl,-rpath," + LIBDIR], check=True, capture_output=True, text=True)
    return exe


def test_c_client_builds_and_fails_loudly_without_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    exe = _build(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 3, r.stdout + r.stderr  # 3 = both constructors returned *_E_NODEVICE
    assert "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_c_client_solves_on_gpu(tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "best_cost=" in r.stdout and "launches=1" in r.stdout


def test_c_latency_client_builds(tmp_path):
    """tools/c_latency.c (bench.py's e2e.c_client: qpmpc_solve timed from plain C) compiles as strict C99 against the
    public header and links against the library alone."""
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    if not os.path.exists(os.path.join(LIBDIR, "libqpmpc_b200.so")):
        import __graft_entry__
        __graft_entry__.build()
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tools", "c_latency.c"), "-o", str(tmp_path / "c_latency"), "-L", LIBDIR,
                    "-lqpmpc_b200", "-lm", "-Wl,-rpath," + LIBDIR], check=True, capture_output=True, text=True)
