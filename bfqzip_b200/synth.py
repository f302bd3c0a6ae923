"""Seeded synthetic read collections.

The reference ships no generator: ``randomFASTQ.py`` only reshuffles an existing FASTQ with an
unseeded ``random.shuffle`` (randomFASTQ.py:82-84).  This module produces the
workloads BASELINE.json names: a random genome with a second haplotype (one SNP every
``snp_every`` bp), reads drawn uniformly from a random haplotype on the forward strand,
``err`` substitution errors per base with low qualities on the error bases, and an ``N``
with quality ``#`` at position 7 of every ``n_every``-th read (the reference needs at least
one ``N`` in the collection).

Reads are returned in the flat layout the C ABI takes: concatenated bases, concatenated
qualities and ``nreads + 1`` uint64 offsets.
"""
from __future__ import annotations

import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


def _mutate(base_codes: np.ndarray, rng: np.random.Generator) -> np.ndarray:
    """Replace each 0..3 code with a different one."""
    return (base_codes + rng.integers(1, 4, size=base_codes.shape, dtype=np.uint8)) & 3


def make_reads(nreads: int, read_len: int, seed: int, coverage: float = 40.0,
               snp_every: int = 1000, err: float = 0.01, n_every: int = 1000,
               genome_len: int | None = None, var_len: bool = False):
    """Single-end collection.  Returns (bases u8[sum len], quals u8[sum len], off u64[nreads+1])."""
    rng = np.random.default_rng(seed)
    if genome_len is None:
        genome_len = max(read_len + 1, int(nreads * read_len / coverage))
    hap0 = rng.integers(0, 4, size=genome_len, dtype=np.uint8)
    hap1 = hap0.copy()
    snp_pos = np.arange(snp_every // 2, genome_len, snp_every)
    hap1[snp_pos] = _mutate(hap0[snp_pos], rng)
    haps = np.stack([hap0, hap1])

    if var_len:
        lens = rng.integers(max(1, read_len // 2), read_len + 1, size=nreads, dtype=np.int64)
    else:
        lens = np.full(nreads, read_len, dtype=np.int64)
    off = np.zeros(nreads + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:].view(np.int64))
    total = int(off[-1])

    start = rng.integers(0, genome_len - read_len + 1, size=nreads, dtype=np.int64)
    which = rng.integers(0, 2, size=nreads, dtype=np.int64)
    # position of every base inside its read
    rid = np.repeat(np.arange(nreads, dtype=np.int64), lens)
    pos_in_read = np.arange(total, dtype=np.int64) - off[:-1].view(np.int64)[rid]
    codes = haps[which[rid], start[rid] + pos_in_read]

    is_err = rng.random(total) < err
    nerr = int(is_err.sum())
    codes[is_err] = _mutate(codes[is_err], rng)
    quals = rng.integers(25, 42, size=total, dtype=np.uint8)
    quals[is_err] = rng.integers(2, 16, size=nerr, dtype=np.uint8)
    quals += 33
    bases = ACGT[codes]

    # an N with quality '#' at position 7 of every n_every-th read (always read 0)
    nr = np.arange(0, nreads, n_every, dtype=np.int64)
    nr = nr[lens[nr] > 7]
    if nr.size == 0:                       # tiny reads: put it at position 0 of read 0
        bases[0] = ord("N"); quals[0] = ord("#")
    else:
        at = off[:-1].view(np.int64)[nr] + 7
        bases[at] = ord("N"); quals[at] = ord("#")
    return bases, quals, off


def make_paired(npairs: int, read_len: int, seed: int, coverage: float = 40.0,
                fragment: int = 350, snp_every: int = 1000, err: float = 0.01,
                n_every: int = 1000):
    """Paired-end collection: mate 2 is the reverse complement of the fragment end.
    Returns ((bases1, quals1, off1), (bases2, quals2, off2))."""
    rng = np.random.default_rng(seed)
    genome_len = max(fragment + 1, int(2 * npairs * read_len / coverage))
    hap0 = rng.integers(0, 4, size=genome_len, dtype=np.uint8)
    hap1 = hap0.copy()
    snp_pos = np.arange(snp_every // 2, genome_len, snp_every)
    hap1[snp_pos] = _mutate(hap0[snp_pos], rng)
    haps = np.stack([hap0, hap1])
    start = rng.integers(0, genome_len - fragment + 1, size=npairs, dtype=np.int64)
    which = rng.integers(0, 2, size=npairs, dtype=np.int64)
    col = np.arange(read_len, dtype=np.int64)
    out = []
    for mate in (0, 1):
        if mate == 0:
            idx = start[:, None] + col[None, :]
            codes = haps[which[:, None], idx]
        else:
            idx = start[:, None] + (fragment - 1 - col)[None, :]
            codes = 3 - haps[which[:, None], idx]            # complement, read backwards
        codes = codes.reshape(-1)
        total = codes.size
        is_err = rng.random(total) < err
        codes[is_err] = _mutate(codes[is_err], rng)
        quals = rng.integers(25, 42, size=total, dtype=np.uint8)
        quals[is_err] = rng.integers(2, 16, size=int(is_err.sum()), dtype=np.uint8)
        quals += 33
        bases = ACGT[codes]
        nr = np.arange(0, npairs, n_every, dtype=np.int64)
        if read_len > 7:
            bases[nr * read_len + 7] = ord("N"); quals[nr * read_len + 7] = ord("#")
        else:
            bases[0] = ord("N"); quals[0] = ord("#")
        off = (np.arange(npairs + 1, dtype=np.uint64) * np.uint64(read_len))
        out.append((bases, quals, off))
    return tuple(out)
