"""In-process seam by execution: the reference's UNMODIFIED host engine (main.cc,
CuCLARK_hh.hh, file.cc ... compiled from /root/reference/src where they lie) linked to oracle/ref_b200_device.cc
-- `class CuClarkDB<HKMERr>` implemented over the C ABI of libcuclark_b200.so -- must print the CSVs the
reference's own device code printed (tests/golden/ref_csv: host engine + CPU stand-in; tests/golden/ref_gpu:
the reference CUDA kernels on a B200).  The binaries oracle/_ref/cuCLARK_b200, cuCLARK-l_b200 are built by
`make -C oracle ref` (in __graft_entry__.build()) and travel to the GPU box prebuilt."""
import glob
import os
import shutil
import subprocess

import pytest

from test_cli_gpu import _write_ref_db

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REFDIR = os.path.join(ROOT, "oracle", "_ref")
GOLD = os.path.join(HERE, "golden")


def _shim(light):
    p = os.path.join(REFDIR, "cuCLARK-l_b200" if light else "cuCLARK_b200")
    if not os.path.exists(p):
        pytest.skip("oracle/_ref/cuCLARK[-l]_b200 not built (make -C oracle ref needs /root/reference)")
    return p


def _targets(world, d):
    labels = open(os.path.join(world, "labels.txt")).read().split()
    tp = os.path.join(d, "targets.txt")
    with open(tp, "w") as f:
        for i, lab in enumerate(labels):
            f.write(f"{os.path.join(world, 'genomes', f'g{i}.fa')}\t{lab}\n")
    return tp


def _reads(world, d):
    """Copies of the read files: the reference's paired mode writes <file1>_ConcatenatedByCLARK.fa beside file1."""
    out = os.path.join(d, "reads")
    os.makedirs(out, exist_ok=True)
    for f in ("reads.fa", "reads.fq", "reads_ml.fa", "p1.fq", "p2.fq"):
        shutil.copy(os.path.join(world, f), out)
    return out


RUNS = {"fa": ["-O", "reads.fa"], "ml": ["-O", "reads_ml.fa"], "fq": ["-O", "reads.fq"], "pe": ["-P", "p1.fq", "p2.fq"],
        "fa_ext": ["-O", "reads.fa", "--extended"], "pe_ext": ["-P", "p1.fq", "p2.fq", "--extended"],
        "fa_b4": ["-O", "reads.fa", "-b", "4"]}


def _run(exe, base, case, rd, res, cwd):
    args = [a if a.startswith("-") or a.isdigit() else os.path.join(rd, a) for a in RUNS[case]]
    r = subprocess.run([exe] + base + args + ["-R", res], cwd=cwd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (case, r.stdout[-800:], r.stderr[-1500:])
    return r


@pytest.mark.parametrize("mode", ["light", "full"])
def test_reference_host_engine_over_the_c_abi_equals_reference_csvs(tmp_path, mode):
    """Databases written by the unmodified reference (tests/golden/ref_db), reads of tests/golden/world: every CSV in
    tests/golden/ref_csv for this mode, byte for byte."""
    light = mode == "light"
    exe = _shim(light)
    d = str(tmp_path)
    world = os.path.join(GOLD, "world")
    tp = _targets(world, d)
    os.makedirs(os.path.join(d, "db"))
    _write_ref_db(os.path.join(GOLD, "ref_db", mode + ".npz"), os.path.join(d, "db"))
    rd = _reads(world, d)
    base = ["-k", "27" if light else "31", "-T", tp, "-D", os.path.join(d, "db") + "/"]
    golden = sorted(glob.glob(os.path.join(GOLD, "ref_csv", f"{mode}_*.csv")))
    assert golden
    for g in golden:
        case = os.path.basename(g)[len(mode) + 1:-4]
        res = os.path.join(d, "out_" + case)
        r = _run(exe, base, case, rd, res, d)
        assert open(res + ".csv", "rb").read() == open(g, "rb").read(), (case, r.stderr[-600:])
    assert not glob.glob(os.path.join(rd, "*_Concatenated*"))      # the reference removes its temp file itself


@pytest.mark.parametrize("mode", ["light", "full", "fullk21"])
def test_reference_host_engine_over_the_c_abi_equals_reference_gpu_kernels(tmp_path, mode):
    """world32 (32 targets): the CSVs the reference's own CUDA kernels wrote on a B200 (tests/golden/ref_gpu/w32_*).
    The database files are written by bin/cuCLARK[-l] first (byte-identical to the reference's, test_build_gpu.py),
    as make_golden.py gpu32 did, so that the shim binary only loads them."""
    light = mode == "light"
    exe = _shim(light)
    d = str(tmp_path)
    world = os.path.join(GOLD, "world32")
    tp = _targets(world, d)
    os.makedirs(os.path.join(d, "db"))
    rd = _reads(world, d)
    k = {"light": "27", "full": "31", "fullk21": "21"}[mode]
    base = ["-k", k, "-T", tp, "-D", os.path.join(d, "db") + "/"]
    mine = os.path.join(ROOT, "bin", "cuCLARK-l" if light else "cuCLARK")
    r = subprocess.run([mine] + base + ["-O", os.path.join(rd, "reads.fa"), "-R", os.path.join(d, "mine")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-1500:]
    golden = sorted(glob.glob(os.path.join(GOLD, "ref_gpu", f"w32_{mode}_*.csv")))
    assert golden
    for g in golden:
        case = os.path.basename(g)[len(f"w32_{mode}_"):-4]
        res = os.path.join(d, "out_" + case)
        r = _run(exe, base, case, rd, res, d)
        assert open(res + ".csv", "rb").read() == open(g, "rb").read(), (case, r.stderr[-600:])
