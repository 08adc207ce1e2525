"""Process boundary through its real callers: the reference's own
scripts/classify_metagenome.sh (staged verbatim as oracle/_ref/scripts/classify_metagenome.sh by `make -C oracle ref`)
and its coordinator `kent -c` (app/kent.cpp compiled as oracle/_ref/kent) drive bin/cuCLARK and bin/cuCLARK-l of this
repository, unchanged: a tree  <root>/bin/{cuCLARK,cuCLARK-l}, <root>/scripts/{classify_metagenome.sh,.settings}
as the reference lays it out (scripts/set_targets.sh:131-146 writes `-T ...` and `-D ...` into .settings,
classify_metagenome.sh:84-158 appends them and calls ../bin/cuCLARK[-l]).  CSVs must equal the goldens the reference's
own engine wrote (tests/golden/ref_csv)."""
import gzip
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
WORLD = os.path.join(GOLD, "world")


@pytest.fixture(scope="module")
def tree(tmp_path_factory):
    script = os.path.join(REFDIR, "scripts", "classify_metagenome.sh")
    if not os.path.exists(script):
        pytest.skip("oracle/_ref/scripts/classify_metagenome.sh not staged (make -C oracle ref needs /root/reference)")
    root = str(tmp_path_factory.mktemp("install"))
    os.makedirs(os.path.join(root, "bin")); os.makedirs(os.path.join(root, "scripts")); os.makedirs(os.path.join(root, "results"))
    for b in ("cuCLARK", "cuCLARK-l"):
        os.symlink(os.path.join(ROOT, "bin", b), os.path.join(root, "bin", b))
    shutil.copy(script, os.path.join(root, "scripts", "classify_metagenome.sh"))
    assert open(os.path.join(root, "scripts", "classify_metagenome.sh"), "rb").read() == open(script, "rb").read()
    labels = open(os.path.join(WORLD, "labels.txt")).read().split()
    tp = os.path.join(root, "targets.txt")
    with open(tp, "w") as f:
        for i, lab in enumerate(labels):
            f.write(f"{os.path.join(WORLD, 'genomes', f'g{i}.fa')}\t{lab}\n")
    # one database directory holds the light and the full files (their names differ), as set_targets.sh sets it up
    db = os.path.join(root, "DB", "custom_0")
    os.makedirs(db)
    for mode in ("light", "full"):
        _write_ref_db(os.path.join(GOLD, "ref_db", mode + ".npz"), db)
    with open(os.path.join(root, "scripts", ".settings"), "w") as f:
        f.write(f"-T {tp}\n-D {db}/\n")
    reads = os.path.join(root, "reads")
    os.makedirs(reads)
    for f in ("reads.fa", "reads.fq", "p1.fq", "p2.fq"):
        shutil.copy(os.path.join(WORLD, f), reads)
    return {"root": root, "reads": reads}


def _script(tree, args):
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, "jn_cuclark_b200") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run(["./classify_metagenome.sh"] + args, cwd=os.path.join(tree["root"], "scripts"), capture_output=True, text=True,
                       timeout=600, env=env)
    assert r.returncode == 0, (r.stdout[-800:], r.stderr[-1500:])
    return r


def _same(path, golden):
    assert open(path, "rb").read() == open(os.path.join(GOLD, "ref_csv", golden), "rb").read(), golden


def test_classify_metagenome_sh_light(tree, tmp_path):
    rd = tree["reads"]
    _script(tree, ["-O", os.path.join(rd, "reads.fa"), "-R", str(tmp_path / "fa"), "--light"]); _same(str(tmp_path / "fa.csv"), "light_fa.csv")
    _script(tree, ["-O", os.path.join(rd, "reads.fq"), "-R", str(tmp_path / "fq"), "--light", "-n", "2"]); _same(str(tmp_path / "fq.csv"), "light_fq.csv")
    _script(tree, ["-P", os.path.join(rd, "p1.fq"), os.path.join(rd, "p2.fq"), "-R", str(tmp_path / "pe"), "--light"]); _same(str(tmp_path / "pe.csv"), "light_pe.csv")
    _script(tree, ["--light", "-O", os.path.join(rd, "reads.fa"), "-R", str(tmp_path / "ext"), "--extended"]); _same(str(tmp_path / "ext.csv"), "light_fa_ext.csv")
    _script(tree, ["-P", os.path.join(rd, "p1.fq"), os.path.join(rd, "p2.fq"), "-R", str(tmp_path / "pex"), "--light", "--extended", "-b", "4"])
    _same(str(tmp_path / "pex.csv"), "light_pe_ext.csv")


def test_classify_metagenome_sh_full(tree, tmp_path):
    rd = tree["reads"]
    _script(tree, ["-k", "31", "-O", os.path.join(rd, "reads.fq"), "-R", str(tmp_path / "fq")]); _same(str(tmp_path / "fq.csv"), "full_fq.csv")
    _script(tree, ["-O", os.path.join(rd, "reads.fa"), "-R", str(tmp_path / "ext"), "--extended", "-n", "3"]); _same(str(tmp_path / "ext.csv"), "full_fa_ext.csv")
    _script(tree, ["-P", os.path.join(rd, "p1.fq"), os.path.join(rd, "p2.fq"), "-R", str(tmp_path / "pex"), "--extended"])
    _same(str(tmp_path / "pex.csv"), "full_pe_ext.csv")


def test_classify_metagenome_sh_gzipped(tree, tmp_path):
    """--gzipped: the script copies the file to a dot-file beside it, gunzips that, classifies, removes it
    (classify_metagenome.sh:118-137, 160-168); file names are relative to scripts/."""
    sd = os.path.join(tree["root"], "scripts")
    for f in ("reads.fq", "p1.fq", "p2.fq"):
        with open(os.path.join(tree["reads"], f), "rb") as src, gzip.open(os.path.join(sd, f + ".gz"), "wb") as dst:
            dst.write(src.read())
    _script(tree, ["-O", "reads.fq.gz", "-R", str(tmp_path / "fq"), "--light", "--gzipped"]); _same(str(tmp_path / "fq.csv"), "light_fq.csv")
    _script(tree, ["-P", "p1.fq.gz", "p2.fq.gz", "-R", str(tmp_path / "pe"), "--gzipped", "--light"]); _same(str(tmp_path / "pe.csv"), "light_pe.csv")
    assert not [f for f in os.listdir(sd) if f.startswith(".reads") or f.startswith(".p1") or f.startswith(".p2")]


def test_kent_c(tree, tmp_path):
    """`kent -c -O reads -R result` (app/kent.cpp:512-545): cd scripts && ./classify_metagenome.sh ... -b 32 --light."""
    kent = os.path.join(REFDIR, "kent")
    if not os.path.exists(kent):
        pytest.skip("oracle/_ref/kent not built")
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, "jn_cuclark_b200") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    rd = tree["reads"]

    def run(args):
        r = subprocess.run([kent, "-c"] + args, cwd=tree["root"], capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0, (r.stdout[-800:], r.stderr[-1500:])
        return r

    run(["-O", os.path.join(rd, "reads.fq"), "-R", str(tmp_path / "k1")]); _same(str(tmp_path / "k1.csv"), "light_fq.csv")
    run(["-O", os.path.join(rd, "reads.fa"), "-R", "k2", "-n", "2", "--extended"])           # relative -R lands in <root>/results/
    _same(os.path.join(tree["root"], "results", "k2.csv"), "light_fa_ext.csv")
    run(["-P", os.path.join(rd, "p1.fq"), os.path.join(rd, "p2.fq"), "-R", str(tmp_path / "k3")]); _same(str(tmp_path / "k3.csv"), "light_pe.csv")
