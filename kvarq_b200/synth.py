"""Deterministic synthetic workloads of the shapes BASELINE.json names.

The reference ships no 4.4 Mb ancestor genome (`.MISSING_LARGE_BLOBS`), so configs 2-5 run on
an H37Rv-like stand-in: an i.i.d. genome with GC = 65.6 %, an MTBC-shaped template set
(27 + 58 SNP templates of 51 bp, regions of 53/131/113/611 bp, 43 spacers of 25 bp; the same
counts and lengths as testsuites/MTBC/{phylo,resistance,spoligo}.py at spacing 25) and
Illumina-profile reads.  Everything is numpy on the host; nothing here is timed.
"""
import numpy as np

GENOME_LEN = 4411532            # testsuites/MTBC/_util.py:9 asserts size 4411533 incl. '\n'
GENOME_SEED = 20141009
SPACING = 25                    # kvarq/config.py:10

_COMP = np.zeros(256, dtype=np.uint8)
for _a, _b in zip(b'ACGTN', b'TGCAN'):
    _COMP[_a] = _b
_BASES = np.frombuffer(b'ACGT', dtype=np.uint8)


def revcomp(seq):
    """Sequence.reverse (kvarq/genes.py:257-262) for bytes over ACGTN."""
    a = np.frombuffer(seq, dtype=np.uint8) if isinstance(seq, (bytes, bytearray)) else seq
    return _COMP[a[::-1]].tobytes()


def genome(length=GENOME_LEN, seed=GENOME_SEED, gc=0.656):
    rng = np.random.default_rng(seed)
    p = np.array([(1 - gc) / 2, gc / 2, gc / 2, (1 - gc) / 2])      # A C G T
    return _BASES[rng.choice(4, size=length, p=p)].copy()


class TemplateSet(object):
    """plus-strand templates + their genome coordinates; .seqs() = plus + minus like
    kvarq/analyse.py:352-354."""

    def __init__(self, plus, positions, kinds):
        self.plus = plus                # list of bytes
        self.positions = positions      # genome start of every template (-1: not on the genome)
        self.kinds = kinds

    @property
    def nplus(self):
        return len(self.plus)

    def seqs(self):
        return list(self.plus) + [revcomp(s) for s in self.plus]


def mtbc_templates(g, seed=1, n_phylo=27, n_res=58, regions=(3, 81, 63, 561), n_spacers=43,
                   spacing=SPACING):
    """MTBC-shaped set on genome g: SNP templates carry the *mutant* base at their centre
    (SNP.seq, kvarq/genes.py:498-501); regions and spacers are copied from the genome."""
    rng = np.random.default_rng(seed)
    plus, pos, kinds = [], [], []
    n_snp = n_phylo + n_res
    centres = np.sort(rng.choice(np.arange(1000, len(g) - 1000), size=n_snp, replace=False))
    for i, c in enumerate(centres):
        t = g[c - spacing:c + spacing + 1].copy()
        alts = [b for b in b'ACGT' if b != t[spacing]]
        t[spacing] = alts[int(rng.integers(0, 3))]
        plus.append(t.tobytes()); pos.append(int(c - spacing)); kinds.append('phylo' if i < n_phylo else 'res')
    for ln in regions:
        s = int(rng.integers(1000, len(g) - 2000))
        plus.append(g[s - spacing:s + ln + spacing].tobytes()); pos.append(s - spacing); kinds.append('region')
    dr = int(rng.integers(1000, len(g) - 10000))
    for i in range(n_spacers):
        s = dr + i * 72                 # spacers sit ~36 bp apart between direct repeats
        plus.append(g[s:s + 25].tobytes()); pos.append(s); kinds.append('spacer')
    return TemplateSet(plus, pos, kinds)


def snp_panel(g, n, seed=4, spacing=SPACING):
    """n random 51-bp SNP templates (config 4: thousands of templates)."""
    rng = np.random.default_rng(seed)
    centres = np.sort(rng.choice(np.arange(1000, len(g) - 1000), size=n, replace=False))
    plus, pos = [], []
    for c in centres:
        t = g[c - spacing:c + spacing + 1].copy()
        alts = [b for b in b'ACGT' if b != t[spacing]]
        t[spacing] = alts[int(rng.integers(0, 3))]
        plus.append(t.tobytes()); pos.append(int(c - spacing))
    return TemplateSet(plus, pos, ['panel'] * n)


def isolate(g, templates, fraction=0.3, seed=7):
    """Isolate genome: a fraction of the SNP templates' mutant bases applied to g."""
    rng = np.random.default_rng(seed)
    h = g.copy()
    for t, p, k in zip(templates.plus, templates.positions, templates.kinds):
        if k in ('phylo', 'res', 'panel') and rng.random() < fraction:
            h[p + SPACING] = t[SPACING]
    return h


def _digits(n):
    return len(str(n))


def fastq_reads(g, nreads, readlen=150, seed=2, qmean0=38.0, qdrop=12.0, qsd=3.0, n_rate=0.001,
                first_index=0, mate=1, block=200000, starts=None, strands=None):
    """FastQ bytes (numpy uint8) of `nreads` reads: uniform start, strand 50/50, header
    '@SYN.<i> <i>/<mate>', Phred+33 with Q(p) = clip(round(N(qmean0 - qdrop (p/L)^2, qsd)), 2, 41),
    substitutions with probability 10^(-Q/10), `n_rate` N calls at Q = 2."""
    rng = np.random.default_rng(seed)
    L = readlen
    qm = (qmean0 - qdrop * (np.arange(L) / float(L)) ** 2).astype(np.float32)
    out = []
    done = 0
    while done < nreads:
        # keep the header width constant inside a block
        i0 = first_index + done
        lim = 10 ** _digits(max(i0, 1)) if i0 > 0 else 10
        nb = int(min(block, nreads - done, lim - i0))
        d = _digits(i0)
        st = rng.integers(0, len(g) - L + 1, size=nb) if starts is None else starts[done:done + nb]
        sd = rng.random(nb) < 0.5 if strands is None else strands[done:done + nb]
        bases = g[st[:, None] + np.arange(L)[None, :]]
        if sd.any():
            bases[sd] = _COMP[bases[sd][:, ::-1]]
        q = np.rint(rng.standard_normal((nb, L), dtype=np.float32) * np.float32(qsd) + qm[None, :])
        q = np.clip(q, 2, 41).astype(np.uint8)
        perr = np.float32(10.0) ** (-q.astype(np.float32) / np.float32(10.0))
        err = rng.random((nb, L), dtype=np.float32) < perr
        if err.any():
            code = (bases[err] >> 1) & 3                    # A0 C1 T2 G3
            alt = np.frombuffer(b'ACTG', dtype=np.uint8)[(code + rng.integers(1, 4, size=code.size)) & 3]
            bases[err] = alt
        isn = rng.random((nb, L), dtype=np.float32) < n_rate
        bases[isn] = ord('N'); q[isn] = 2
        W = 5 + d + 1 + d + 3 + L + 3 + L + 1
        rec = np.empty((nb, W), dtype=np.uint8)
        idx = np.arange(i0, i0 + nb)
        dig = np.empty((nb, d), dtype=np.uint8)
        for k in range(d):
            dig[:, d - 1 - k] = (idx // 10 ** k) % 10 + 48
        c = 0
        rec[:, c:c + 5] = np.frombuffer(b'@SYN.', dtype=np.uint8); c += 5
        rec[:, c:c + d] = dig; c += d
        rec[:, c] = 32; c += 1
        rec[:, c:c + d] = dig; c += d
        rec[:, c:c + 3] = np.frombuffer(b'/%d\n' % mate, dtype=np.uint8); c += 3
        rec[:, c:c + L] = bases; c += L
        rec[:, c:c + 3] = np.frombuffer(b'\n+\n', dtype=np.uint8); c += 3
        rec[:, c:c + L] = q + 33; c += L
        rec[:, c] = 10
        out.append(rec.reshape(-1))
        done += nb
    return np.concatenate(out) if len(out) != 1 else out[0]


def count_bases(fastq):
    """Untrimmed bases = total length of the sequence lines (the Gbp numerator)."""
    a = np.frombuffer(fastq, dtype=np.uint8) if not isinstance(fastq, np.ndarray) else fastq
    nl = np.flatnonzero(a == 10)
    n = (len(nl) // 4) * 4
    nl = nl[:n].reshape(-1, 4)
    return int((nl[:, 1] - nl[:, 0] - 1).sum())


# ------------------------------------------------------ the MTBC suites at their real positions
def mtbc_genome(seed=GENOME_SEED):
    """Synthetic stand-in for the MTB ancestor genome the reference does not ship:
    the i.i.d. genome with the ORIGINAL base forced at every resistance-SNP position
    (testsuites/MTBC/resistance.py:136-193) and a base other than the mutant at every phylogenetic
    SNP position (phylo.py:132-162), so that the assertions of SNP.__init__ (kvarq/genes.py:491-495)
    hold.  Positions are 1-based."""
    from . import testsuite_data as D
    g = genome(seed=seed)
    for pos, orig, _, _ in D.RESISTANCE_SNPS:
        g[pos - 1] = ord(orig)
    for pos, base, _ in D.PHYLO_SNPS:
        if g[pos - 1] == ord(base):
            g[pos - 1] = [b for b in b'ACGT' if b != ord(base)][pos % 3]
    return g


def plant(g, mutations):
    """isolate = genome with (1-based position, base) substitutions"""
    h = g.copy()
    for pos, base in mutations:
        h[pos - 1] = ord(base) if isinstance(base, str) else base
    return h


def reads_around(g, sites, per_site=60, readlen=150, seed=3, **kw):
    """FastQ of reads that overlap the given (1-based start, length) sites of genome g: `per_site` reads per
    site, starts uniform over every position where the read still touches the site."""
    rng = np.random.default_rng(seed)
    starts = []
    for pos, length in sites:
        lo, hi = max(0, pos - 1 - readlen + 8), min(len(g) - readlen, pos - 1 + length - 8)
        starts.append(rng.integers(lo, max(lo + 1, hi), size=per_site))
    starts = np.concatenate(starts)
    rng.shuffle(starts)
    return fastq_reads(g, len(starts), readlen=readlen, seed=seed + 1, starts=starts, **kw)
