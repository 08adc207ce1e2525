"""Deterministic synthetic genomes / reads for the parity tests (numpy only).

The generator deliberately produces the quirks the reference's parser and
database builder react to: lower-case
bases, N runs, multi-record FASTA targets, segments shared between targets
(non-discriminative k-mers), two files under one label, reads shorter than k,
reads with N at the start/middle/end, multi-line FASTA reads, long names.
"""
from __future__ import annotations

import os
import numpy as np

BASES = np.frombuffer(b"ACGT", dtype=np.uint8)
COMP = {65: 84, 67: 71, 71: 67, 84: 65, 78: 78}


def rand_seq(rng: np.random.Generator, n: int) -> np.ndarray:
    return BASES[rng.integers(0, 4, size=n)]


def revcomp(s: np.ndarray) -> np.ndarray:
    lut = np.zeros(256, dtype=np.uint8)
    for a, b in COMP.items():
        lut[a] = b
        lut[a + 32] = b + 32 if a != 78 else 110
    return lut[s[::-1]]


def write_fasta(path: str, records: list[tuple[str, np.ndarray]], width: int = 80) -> None:
    with open(path, "wb") as f:
        for name, seq in records:
            f.write(b">" + name.encode() + b"\n")
            b = seq.tobytes()
            for i in range(0, len(b), width):
                f.write(b[i:i + width] + b"\n")


def make_genomes(out_dir: str, n_genomes: int = 6, length: int = 20000, seed: int = 1,
                 shared: bool = True, quirks: bool = True, same_label_pair: bool = True):
    """Writes <out_dir>/g<i>.fa and targets.txt; returns (targets_path, genomes, labels)."""
    os.makedirs(out_dir, exist_ok=True)
    rng = np.random.default_rng(seed)
    genomes = [rand_seq(rng, length) for _ in range(n_genomes)]
    if shared and n_genomes >= 3:
        # a 600-base block shared by g0 and g1 (and reverse-complemented into g2)
        blk = rand_seq(rng, 600)
        genomes[0][1000:1600] = blk
        genomes[1][5000:5600] = blk
        genomes[2][7000:7600] = revcomp(blk)
    if shared and n_genomes >= 6:
        # same offsets => same window phase in light mode (non-overlapping k-mers)
        genomes[5][:3000] = genomes[3][:3000]             # different labels -> removed
        genomes[5][10000:12000] = genomes[4][10000:12000] # same label (T4) -> kept
    if quirks:
        genomes[0][300:310] = ord("N")              # N run
        genomes[1][40:45] |= 0x20                   # lower case acgt
        genomes[min(3, n_genomes - 1)][min(9000, length - 1)] = ord("R")   # IUPAC code
    labels = []
    lines = []
    for i, g in enumerate(genomes):
        p = os.path.join(out_dir, f"g{i}.fa")
        if quirks and i == 1:
            # two records in one file
            write_fasta(p, [(f"g{i}_a some description", g[: length // 2]), (f"g{i}_b", g[length // 2:])], 70)
        else:
            write_fasta(p, [(f"g{i}", g)], 80 if i % 2 == 0 else 61)
        lab = f"T{i}"
        if same_label_pair and i == n_genomes - 1 and n_genomes >= 2:
            lab = f"T{n_genomes - 2}"                # last two files share one label
        labels.append(lab)
        lines.append(f"{os.path.abspath(p)}\t{lab}\n")
    tp = os.path.join(out_dir, "targets.txt")
    with open(tp, "w") as f:
        f.writelines(lines)
    return tp, genomes, labels


def make_reads(genomes: list[np.ndarray], n: int, length: int = 150, seed: int = 2,
               err: float = 0.01, edge_cases: bool = True, k: int = 31):
    """Returns list of (name, seq bytes ndarray)."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        u = rng.random()
        if u < 0.85:
            g = genomes[rng.integers(0, len(genomes))]
            p = int(rng.integers(0, len(g) - length + 1))
            s = g[p:p + length].copy()
            if rng.random() < 0.5:
                s = revcomp(s)
            m = rng.random(length) < err
            s[m] = BASES[rng.integers(0, 4, size=int(m.sum()))]
        elif u < 0.95:
            s = rand_seq(rng, length)
        else:
            # chimeric read: half from one genome, half from another (two targets hit)
            g1 = genomes[rng.integers(0, len(genomes))]
            g2 = genomes[rng.integers(0, len(genomes))]
            h = length // 2
            p1 = int(rng.integers(0, len(g1) - h)); p2 = int(rng.integers(0, len(g2) - (length - h)))
            s = np.concatenate([g1[p1:p1 + h], g2[p2:p2 + length - h]])
        if rng.random() < 0.02:
            s = s.copy(); s[int(rng.integers(0, length))] = ord("N")
        out.append((f"r{i}", s))
    if edge_cases:
        g = genomes[0]
        base = g[2000:2000 + length].copy()
        def mod(fn):
            s = base.copy(); fn(s); return s
        extra = [
            ("short_k_minus_1", g[2000:2000 + k - 1].copy()),
            ("exact_k", g[2000:2000 + k].copy()),
            ("k_plus_1", g[2000:2000 + k + 1].copy()),
            ("tiny", g[2000:2003].copy()),
            ("n_first", mod(lambda s: s.__setitem__(0, ord("N")))),
            ("n_last", mod(lambda s: s.__setitem__(length - 1, ord("N")))),
            ("n_mid", mod(lambda s: s.__setitem__(length // 2, ord("N")))),
            ("n_every_20", mod(lambda s: s.__setitem__(slice(10, None, 20), ord("N")))),
            ("all_n", np.full(length, ord("N"), dtype=np.uint8)),
            ("lower_case", base | 0x20),
            ("rna_u", mod(lambda s: s.__setitem__(s == ord("T"), ord("U")))),
            ("short_part_then_long", np.concatenate([g[2000:2010], [ord("N")], g[2011:2100]]).astype(np.uint8)),
            ("long_then_short_part", np.concatenate([g[2000:2090], [ord("-")], g[2091:2100]]).astype(np.uint8)),
            ("x" * 60 + "_longname", base.copy()),
            ("name with spaces and\ttab", base.copy()),
            ("long_read", g[3000:3000 + 1000].copy()),
        ]
        out.extend(extra)
    return out


def write_reads_fasta(path: str, reads, width: int | None = None) -> None:
    with open(path, "wb") as f:
        for name, s in reads:
            f.write(b">" + name.encode() + b"\n")
            b = s.tobytes()
            if width:
                for i in range(0, max(len(b), 1), width):
                    f.write(b[i:i + width] + b"\n")
            else:
                f.write(b + b"\n")


def write_reads_fastq(path: str, reads, suffix: str = "") -> None:
    with open(path, "wb") as f:
        for name, s in reads:
            f.write(b"@" + name.encode() + suffix.encode() + b"\n" + s.tobytes() + b"\n+\n" + b"I" * len(s) + b"\n")


def make_pairs(genomes, n: int, length: int = 150, insert: int = 400, seed: int = 3, err: float = 0.01):
    rng = np.random.default_rng(seed)
    r1, r2 = [], []
    for i in range(n):
        g = genomes[rng.integers(0, len(genomes))]
        p = int(rng.integers(0, len(g) - insert + 1))
        a = g[p:p + length].copy()
        b = revcomp(g[p + insert - length:p + insert])
        for s in (a, b):
            m = rng.random(length) < err
            s[m] = BASES[rng.integers(0, 4, size=int(m.sum()))]
        if i % 7 == 3:
            b = b[: length - 30]                     # unequal mates
        name = f"p{i}"
        r1.append((name, a)); r2.append((name, b))
    return r1, r2
