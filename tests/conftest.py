import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    import oracle_py
    oracle_py.lib()
    return oracle_py


@pytest.fixture(scope="session")
def small_world(tmp_path_factory):
    """6 genomes x 20 kbp with quirks + 2000 reads (+ edge cases), FASTA/FASTQ/paired files."""
    import synth
    d = str(tmp_path_factory.mktemp("world"))
    tp, genomes, labels = synth.make_genomes(os.path.join(d, "genomes"), 6, 20000, seed=1)
    reads = synth.make_reads(genomes, 2000, 150, seed=2, k=31)
    synth.write_reads_fasta(os.path.join(d, "reads.fa"), reads)
    synth.write_reads_fasta(os.path.join(d, "reads_ml.fa"), reads, width=60)
    synth.write_reads_fastq(os.path.join(d, "reads.fq"), [r for r in reads if " " not in r[0]])
    r1, r2 = synth.make_pairs(genomes, 500)
    synth.write_reads_fastq(os.path.join(d, "p1.fq"), r1, "/1")
    synth.write_reads_fastq(os.path.join(d, "p2.fq"), r2, "/2")
    return {"dir": d, "targets": tp, "genomes": genomes, "labels": labels}
