"""CPU-side checks of the C-ABI boundary: the library loads and exports every declared symbol."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header="cuclark_b200.h"):
    h = open(os.path.join(ROOT, "include", header)).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(ccb_[a-z0-9_]+)\s*\(", h)))


def test_host_header_symbols_exported_and_bound():
    from jn_cuclark_b200 import host, lib
    L = lib.load_library()
    names = _declared("ccb_host.h")
    assert len(names) >= 10
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/ccb_host.h but not exported"
        assert n in host.HOST_SYMBOLS, f"{n} has no ctypes prototype"


def test_cli_binaries_built():
    import subprocess
    for b in ("cuCLARK", "cuCLARK-l"):
        p = os.path.join(ROOT, "bin", b)
        assert os.path.exists(p), p + " missing: run make -C jn_cuclark_b200"
        r = subprocess.run([p, "--version"], capture_output=True, text=True)
        assert r.returncode == 0 and "sm_100a" in r.stdout
        r = subprocess.run([p, "-k", "31", "-T", "/nonexistent", "-D", "/tmp", "-O", "x", "-R", "y"], capture_output=True, text=True)
        assert r.returncode == 1 and "Failed to find/read the file of the targets definition" in r.stderr
        r = subprocess.run([p, "-k", "31", "-Q", "zz", "-D", "/tmp", "-O", "x"], capture_output=True, text=True)
        assert r.returncode == 1 and "Failed to recognize option: -Q" in r.stderr


def test_header_symbols_exported_and_bound():
    from jn_cuclark_b200 import lib
    L = lib.load_library()
    names = _declared()
    assert len(names) >= 19
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/cuclark_b200.h but not exported"
        assert n in lib.SYMBOLS, f"{n} has no ctypes prototype"
    assert L.ccb_version().decode().endswith("sm_100a")


def test_no_gpu_fails_loudly():
    """Without a CUDA device the product refuses to run (no CPU fallback)."""
    import torch
    import pytest
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from jn_cuclark_b200 import CuClarkDB, CcbError
    with pytest.raises(CcbError) as e:
        CuClarkDB(1, 31, 1, 4)
    assert "No CUDA devices found" in str(e.value)


def test_product_does_not_reference_oracle():
    bad = []
    for d, _, fs in os.walk(os.path.join(ROOT, "jn_cuclark_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp", "Makefile")):
                s = open(os.path.join(d, f), errors="ignore").read()
                if re.search(r"oracle|ccbo_", s):
                    bad.append(os.path.join(d, f))
    assert not bad, bad


def test_flag_values_match_the_header():
    """The query flags of include/cuclark_b200.h and their mirror in lib.py are the same numbers."""
    from jn_cuclark_b200 import lib
    h = open(os.path.join(ROOT, "include", "cuclark_b200.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    flags = dict((n, int(v)) for n, v in re.findall(r"\b(CCB_[A-Z_]+)\s*=\s*(\d+)", h))
    assert {"CCB_EXTENDED", "CCB_STATS", "CCB_FORCE_DENSE", "CCB_NO_ANCHOR", "CCB_WARP_KERNEL"} <= set(flags)
    for n, v in flags.items():
        assert getattr(lib, n) == v, n


def test_library_holds_sm_100a_code_only():
    """Every device image in the library is sm_100a (no other architecture, no PTX-only fallback), and a probe is one
    256-bit load with the 64-byte L2 fill hint -- checked on the build, no GPU needed.  (The TMA-staged variant of
    the block kernel is a measurement build now: make VARIANT=exp EXTRA=-DCCB_EXPERIMENTS.)"""
    import shutil
    import subprocess
    import pytest
    from jn_cuclark_b200 import lib
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    r = subprocess.run(["cuobjdump", "-lelf", lib.lib_path()], capture_output=True, text=True)
    elfs = [l for l in r.stdout.splitlines() if l.startswith("ELF file")]
    assert elfs and all("sm_100a" in l for l in elfs), elfs
    sass = subprocess.run(["cuobjdump", "-sass", lib.lib_path()], capture_output=True, text=True).stdout
    assert "ENL2.LTC64B.256" in sass                   # LDG ... .256: one 32-byte table sector per probe, 64-byte fill on a miss
    assert "MATCH.ANY" in sass                         # the one-pass route kernel ranks k-mers per owner with __match_any_sync
