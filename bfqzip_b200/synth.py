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


# ---------------------------------------------------------------------------------------------------------------------
# Counter-based generator (BASELINE configs 4 and 5 in bench.py): every base is a pure function of (seed, read, offset),
# so ANY block of a collection can be generated where it is needed -- on the GPU that processes it -- and the numpy and
# torch versions below give the same bytes (tests/test_parallel.py checks that on small sizes).  Same model as make_reads /
# make_paired: random genome, second haplotype with a SNP every 1000 bp, reads from a random haplotype, 1 % substitution
# errors with low qualities, an N with quality '#' at offset 7 of every 1000th read; mate 2 = reverse complement of the
# fragment end.  All arithmetic is on int64 with wrap-around and logical shifts spelled out, which numpy arrays and torch
# tensors do identically.
_U64 = (1 << 64) - 1


def _s64(c: int) -> int:
    c &= _U64
    return c - (1 << 64) if c >= (1 << 63) else c


_GOLD, _MUL1, _MUL2 = _s64(0x9E3779B97F4A7C15), _s64(0xBF58476D1CE4E5B9), _s64(0x94D049BB133111EB)


def _mix_scalar(x: int) -> int:
    z = (x + 0x9E3779B97F4A7C15) & _U64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _U64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _U64
    return z ^ (z >> 31)


def _mix(x):
    """splitmix64 finaliser on int64 arrays / tensors (two's complement: +, * and ^ wrap; >> is arithmetic, so it is masked)."""
    z = x + _GOLD
    z = (z ^ ((z >> 30) & ((1 << 34) - 1))) * _MUL1
    z = (z ^ ((z >> 27) & ((1 << 37) - 1))) * _MUL2
    return z ^ ((z >> 31) & ((1 << 33) - 1))


def _pos(z):
    """63 non-negative bits of a hash."""
    return (z >> 1) & ((1 << 63) - 1)


def _hap_codes(pos, which, seed):
    """Base code 0..3 of genome position pos on haplotype which (0/1)."""
    k0, k1 = _s64(_mix_scalar(seed * 4 + 0)), _s64(_mix_scalar(seed * 4 + 1))
    c0 = _pos(_mix(pos ^ k0)) & 3
    alt = (c0 + 1 + _pos(_mix(pos ^ k1)) % 3) & 3
    snp = (((pos % 1000) == 500) & (which == 1)) * 1          # 0/1 as integers (a bool cannot be subtracted in torch)
    return c0 * (1 - snp) + alt * snp


def hash_reads_codes(idx, read_len, seed, genome_len, mate, fragment, nreads_total):
    """idx: int64 array/tensor of flat base indices (read * read_len + offset) of ONE mate file.  Returns (byte codes of
    the bases as int64 ASCII values, ASCII qualities) as int64 arrays of the same kind."""
    r = idx // read_len
    j = idx - r * read_len
    k2, k3 = _s64(_mix_scalar(seed * 4 + 2)), _s64(_mix_scalar(seed * 4 + 3))
    hr = _mix(r ^ k2)                                             # per read (pair): start and haplotype, shared by the mates
    start = _pos(hr) % (genome_len - fragment + 1)
    which = (hr & 1)
    if mate == 0:
        code = _hap_codes(start + j, which, seed)
    else:
        code = 3 - _hap_codes(start + (fragment - 1) - j, which, seed)
    u = _pos(_mix((idx + mate * nreads_total * read_len) ^ k3))  # per base: error, replacement, quality
    err = ((u & 0xFFFF) < 655) * 1                                # 1 % substitution errors
    code = (code + err * (1 + ((u >> 16) & 0xFFFF) % 3)) & 3
    qual = 33 + (1 - err) * (25 + ((u >> 32) & 0xFFFF) % 17) + err * (2 + ((u >> 32) & 0xFFFF) % 14)
    # A C G T = 65 67 71 84
    base = 65 + 2 * ((code == 1) * 1) + 6 * ((code == 2) * 1) + 19 * ((code == 3) * 1)
    isn = (((r % 1000) == 0) & (j == min(7, read_len - 1))) * 1
    base = base * (1 - isn) + 78 * isn
    qual = qual * (1 - isn) + 35 * isn
    return base, qual


def hash_genome_len(nreads_all_mates: int, read_len: int, coverage: float, fragment: int) -> int:
    return max(fragment + 1, int(nreads_all_mates * read_len / coverage))


def hash_reads(nreads_total: int, read_len: int, seed: int, genome_len: int, mate: int = 0, fragment: int | None = None,
               lo: int = 0, hi: int | None = None):
    """Reads [lo, hi) of one mate file of the counter-based collection, on the host.  Returns (bases, quals, off)."""
    hi = nreads_total if hi is None else hi
    fragment = read_len if fragment is None else fragment
    idx = np.arange(lo * read_len, hi * read_len, dtype=np.int64)
    b, q = hash_reads_codes(idx, read_len, seed, genome_len, mate, fragment, nreads_total)
    off = np.arange(hi - lo + 1, dtype=np.uint64) * np.uint64(read_len)
    return b.astype(np.uint8), q.astype(np.uint8), off


def hash_paired(npairs: int, read_len: int, seed: int, coverage: float = 40.0, fragment: int = 350):
    """Both mate files of a paired collection (the layout make_paired returns)."""
    g = hash_genome_len(2 * npairs, read_len, coverage, fragment)
    return tuple(hash_reads(npairs, read_len, seed, g, mate, fragment) for mate in (0, 1))


def hash_reads_torch(torch, dev, nreads_total: int, read_len: int, seed: int, genome_len: int, mate: int = 0,
                     fragment: int | None = None, lo: int = 0, hi: int | None = None, out=None, out_at: int = 0,
                     chunk_reads: int = 1 << 19):
    """The same reads generated on the device `dev`, chunk by chunk (int64 temporaries of chunk_reads * read_len elements).
    out = (bases u8, quals u8) tensors to fill from element out_at on, or None to allocate.  Returns (bases, quals)."""
    hi = nreads_total if hi is None else hi
    fragment = read_len if fragment is None else fragment
    n = (hi - lo) * read_len
    if out is None:
        out = (torch.empty(n, dtype=torch.uint8, device=dev), torch.empty(n, dtype=torch.uint8, device=dev))
        out_at = 0
    for a in range(lo, hi, chunk_reads):
        b = min(hi, a + chunk_reads)
        idx = torch.arange(a * read_len, b * read_len, dtype=torch.int64, device=dev)
        cb, cq = hash_reads_codes(idx, read_len, seed, genome_len, mate, fragment, nreads_total)
        s = out_at + (a - lo) * read_len
        out[0][s:s + idx.numel()] = cb.to(torch.uint8)
        out[1][s:s + idx.numel()] = cq.to(torch.uint8)
        del idx, cb, cq
    return out


def hash_paired_block_torch(torch, dev, npairs: int, read_len: int, seed: int, genome_len: int, fragment: int, lo: int, hi: int):
    """One block of a paired collection the way split_fastq_2 (BFQzip_parallel.py:325-360) builds it: mates 1 of pairs
    [lo, hi) followed by their mates 2, generated on the device.  Returns (bases, quals, off as numpy u64)."""
    n = (hi - lo) * read_len
    out = (torch.empty(2 * n, dtype=torch.uint8, device=dev), torch.empty(2 * n, dtype=torch.uint8, device=dev))
    for mate in (0, 1):
        hash_reads_torch(torch, dev, npairs, read_len, seed, genome_len, mate, fragment, lo, hi, out, mate * n)
    off = np.arange(2 * (hi - lo) + 1, dtype=np.uint64) * np.uint64(read_len)
    return out[0], out[1], off
