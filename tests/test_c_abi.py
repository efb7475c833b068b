"""A plain C host on the C ABI (tests/c/abi_driver.c): compiled with gcc against include/femx.h and fem_b200/libfemx.so.
CPU: it must compile and link (every symbol the header declares that the driver uses resolves).  GPU: it runs -- COO arrays
vs CSR pattern, the MPIAIJ diagonal / off-diagonal split with garray, null space and symmetry of K, the enqueue-only
(femx_set_async) form -- the checks a PETSc-side maintainer would make before wiring MatSetPreallocationCOO."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "c", "abi_driver")


def _build():
    lib = os.path.join(ROOT, "fem_b200")
    if not os.path.exists(os.path.join(lib, "libfemx.so")):
        subprocess.check_call(["make", "-C", os.path.join(lib, "csrc"), "-j8"])
    cmd = ["gcc", "-std=c99", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "c", "abi_driver.c"), "-o", EXE,
           "-L", lib, "-lfemx", "-lm", "-Wl,-rpath," + lib, "-Wl,--allow-shlib-undefined"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return EXE


def test_c_driver_compiles_and_links():
    exe = _build()
    assert os.path.exists(exe)
    # without a device the driver must fail loudly at femx_create (no CPU fallback), not crash
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
        assert r.returncode == 2 and "no CUDA device" in r.stderr, r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("n", [2, 7])
def test_c_driver_runs(n):
    exe = _build()
    r = subprocess.run([exe, str(n)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ABI DRIVER OK" in r.stdout


def test_header_library_and_fortran_module_agree():
    """include/femx.h, the exported symbols of libfemx.so, the ctypes table and fortran/femx_mod.F90 name the same entry points."""
    import re
    _build()
    hdr = open(os.path.join(ROOT, "include", "femx.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(femx_[A-Za-z0-9_]+)\s*\(", hdr))
    nm = subprocess.run(["nm", "-D", "--defined-only", os.path.join(ROOT, "fem_b200", "libfemx.so")], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in nm.splitlines() if " T femx_" in l}
    fortran = set(re.findall(r'name="(femx_[A-Za-z0-9_]+)"', open(os.path.join(ROOT, "fortran", "femx_mod.F90")).read()))
    from fem_b200 import _lib
    assert declared == exported, (declared ^ exported)
    assert fortran == exported, (fortran ^ exported)
    assert set(_lib.SIGNATURES) == exported, (set(_lib.SIGNATURES) ^ exported)
