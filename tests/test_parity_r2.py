"""Parity cases added in round 2 (VERDICT r1 "what's weak" 1-4): kernel variants and sizes that no
round-1 test reached.  All through the C ABI, bit-exact against the CPU checkers."""
import os
import random

import numpy as np
import pytest

from util import assert_same_result, coverage_from_hits
from kvarq_b200 import synth

pytestmark = pytest.mark.gpu

DEFAULT = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
COMP = bytes.maketrans(b'ACGT', b'TGCA')


def oracle(ck, files, seqs, cfg, impl=None, nthreads=1):
    impl = impl or ('ref' if ck.available('ref') else 'oracle')
    return ck.findseqs(files, seqs, nthreads=nthreads, impl=impl, **cfg)


def _panel(g, n, length, seed):
    rng = np.random.RandomState(seed)
    plus = [bytes(g[p:p + length]) for p in sorted(rng.choice(len(g) - length - 10, n, replace=False))]
    return plus


# ----------------------------------------------------------------------- negative Amin (NEG) --
@pytest.mark.parametrize('big', [False, True])
@pytest.mark.parametrize('amin', ['\x80', '\xa0', '\xff'])
def test_negative_amin(checker, tmp_path, amin, big):
    """`Amin` >= 0x80 is a negative signed char (workhorse.c:1058 compares `*ptr>=Amin` on plain
    char): every byte 0..127 passes -- the line's '\\n' included, so a run still open at the end
    of the quality line is never closed and does not count (workhorse.c:1055-1068) -- and bytes
    >= 0x80 pass only when they are >= Amin as signed values.  Runs kvq_scan_kernel<true, *>."""
    from kvarq_b200.engine import Engine
    g = synth.genome(60000, seed=71)
    rng = random.Random(72)
    if big:
        plus = _panel(g, 720, 51, 73) + _panel(g, 40, 25, 74)        # > 2048 T1 pieces: large-set variant
    else:
        plus = _panel(g, 60, 51, 73) + _panel(g, 20, 25, 74) + [bytes(g[5000:5131])]
    seqs = [s.decode() for s in plus] + [s.translate(COMP)[::-1].decode() for s in plus]
    qual = [0x85, 0x95, 0x9f, 0xa0, 0xb0, 0xc8, 0xff, 0x49, 0x49, 0x35, 0x7f, 0x21]
    weights = [1, 1, 1, 6, 6, 6, 6, 20, 20, 6, 3, 3]
    recs = []
    for i in range(1500 if big else 2500):
        L = rng.randint(30, 200)
        a = rng.randrange(len(g) - L)
        s = bytes(g[a:a + L])
        if rng.random() < 0.5:
            s = s.translate(COMP)[::-1]
        q = bytes(rng.choices(qual, weights, k=L))
        if i % 11 == 0:
            q = q[:-1] + b'\x85'                 # a bad byte right before the newline closes the last run
        if i % 13 == 0:
            q = bytes([0x49]) * L                # no byte below Amin at all: no run is ever closed, rl = 0
        recs.append(b'@n%d\n%s\n+\n%s\n' % (i, s, q))
    fn = str(tmp_path / 'neg.fastq')
    open(fn, 'wb').write(b''.join(recs))
    cfg = dict(DEFAULT, Amin=amin)
    e = Engine(0)
    e.config(**cfg)
    got = e.findseqs(fn, seqs, nplus=len(plus))
    want = oracle(checker, fn, seqs, cfg)
    assert_same_result(got, want, what='Amin=%r big=%s' % (amin, big))
    if amin == '\xa0':
        assert len(want['hits']) > 40 and len(set(want['stats']['readlengths'])) > 3
    if amin == '\x80':
        assert want['stats']['readlengths'][0] == len(recs)          # nothing is below -128: no run ever closes
    cov, mut = coverage_from_hits(want['hits'], want['hitseqs'], seqs, len(plus))
    assert got.coverage.tolist() == cov and got.mutcount.tolist() == mut
    st = e.stage([np.frombuffer(b''.join(recs), dtype=np.uint8)])
    assert_same_result(e.scan_staged(st, seqs, nplus=len(plus)), want, what='staged Amin=%r big=%s' % (amin, big))
    st.free(); e.close()


def test_maxerrors_beyond_seed_scheme(gpu_engine, checker, tmp_path):
    """engine.config(maxerrors=12): E+1 = 13 pieces is beyond the seeded path (<= 8); the reference
    accepts any maxerrors, so must the engine (exhaustive matcher).  ADVICE r1 engine.cpp:734."""
    rng = random.Random(5)
    rs = lambda n: ''.join(rng.choice('ACGT') for _ in range(n))
    tpl = rs(70)
    recs = []
    for i in range(60):
        L = rng.randint(40, 120)
        s = list(rs(L))
        if i % 2 == 0:
            a = rng.randint(0, L - 30); s[a:a + 70] = list(tpl); s = s[:L]
            for _ in range(rng.randint(0, 14)):
                s[rng.randrange(L)] = rng.choice('ACGT')
        recs.append('@r%d\n%s\n+\n%s\n' % (i, ''.join(s), 'I' * L))
    fn = str(tmp_path / 'e12.fastq')
    open(fn, 'w').write(''.join(recs))
    for E in (12, 9, 23):
        cfg = dict(maxerrors=E, minoverlap=30, minreadlength=30, Amin='5')
        gpu_engine.config(**cfg)
        got = gpu_engine.findseqs(fn, [tpl, synth.revcomp(tpl.encode()).decode()])
        want = oracle(checker, fn, [tpl, synth.revcomp(tpl.encode()).decode()], cfg)
        assert len(want['hits']) > 20
        assert_same_result(got, want, what='maxerrors=%d' % E)


def test_readlength_histogram_beyond_1024(gpu_engine, checker, tmp_path):
    """add_rl (workhorse.c:394-402): rl >= 1024 is not counted but still raises rls_longest; the
    stats tuple stops at 1024 entries (the reference would read past rls_buf there)."""
    rng = random.Random(15)
    rs = lambda n: ''.join(rng.choice('ACGT') for _ in range(n))
    tpl = rs(60)
    recs = []
    for i, L in enumerate([100, 1023, 1024, 1025, 2000, 150, 5000, 1023]):
        s = list(rs(L)); s[10:70] = list(tpl)
        recs.append('@r%d\n%s\n+\n%s\n' % (i, ''.join(s), 'I' * L))
    fn = str(tmp_path / 'rl1024.fastq')
    open(fn, 'w').write(''.join(recs))
    cfg = dict(maxerrors=1, minoverlap=20, minreadlength=20, Amin='5')
    gpu_engine.config(**cfg)
    got = gpu_engine.findseqs(fn, [tpl])
    want = oracle(checker, fn, [tpl], cfg)
    assert len(want['stats']['readlengths']) == 1024 and want['stats']['readlengths'][1023] == 2
    assert_same_result(got, want, what='rl >= 1024')


# -------------------------------------------------------------- config-2 shape at real size ---
def test_config2_shape_200k_reads(checker, tmp_path):
    """BASELINE config 2 at 200 000 reads x the 264 MTBC-shaped seqs: stats, per-template counters
    and fused coverage against the multi-threaded restatement (deterministic: oracle/kvarq_oracle.c
    cuts chunks like nthreads=1), hits bit-exact in order; and the first 20 000 reads against the
    reference's own C (oracle/_ref, nthreads=1)."""
    from kvarq_b200.engine import Engine
    g = synth.genome()
    ts = synth.mtbc_templates(g, seed=1)
    iso = synth.isolate(g, ts, seed=7)
    fq = synth.fastq_reads(iso, 200000, readlen=150, seed=2)
    fn = str(tmp_path / 'c2_200k.fastq')
    open(fn, 'wb').write(fq.tobytes())
    seqs = ts.seqs()
    e = Engine(0)
    e.config(**DEFAULT)
    got = e.findseqs(fn, seqs, nplus=ts.nplus)
    want = oracle(checker, fn, seqs, DEFAULT, impl='oracle', nthreads=os.cpu_count() or 8)
    assert want['stats']['records_parsed'] == 200000 and len(want['hits']) > 500
    assert_same_result(got, want, what='config 2, 200k reads vs oracle (threads)')
    cov, mut = coverage_from_hits(want['hits'], want['hitseqs'], seqs, ts.nplus)
    assert got.coverage.tolist() == cov and got.mutcount.tolist() == mut
    st = e.stage([fq])
    res = e.scan_staged(st, seqs, nplus=ts.nplus)
    assert_same_result(res, want, what='config 2, 200k reads, resident')
    assert [tuple(h) for h in res['hits']] == [tuple(h) for h in got['hits']]
    st.free()
    if checker.available('ref'):
        nl = np.flatnonzero(fq[:20000 * 400] == 10)
        pre = fq[:int(nl[4 * 20000 - 1]) + 1].tobytes()
        fn2 = str(tmp_path / 'c2_20k.fastq')
        open(fn2, 'wb').write(pre)
        got2 = e.findseqs(fn2, seqs, nplus=ts.nplus)
        want2 = oracle(checker, fn2, seqs, DEFAULT, impl='ref', nthreads=1)
        assert_same_result(got2, want2, what='config 2, 20k-read prefix vs the reference C')
    e.close()


# ------------------------------------------------------------- chunk-parallel inflate (N4) ------
def test_large_gz_parallel_inflate(checker, tmp_path, monkeypatch):
    """.gz inputs large enough for parallel inflate rounds (kvarq_b200/csrc/gz_parallel.h): single
    member and three members + trailing junk, two files side by side.  Hits, hitseqs and stats -- the
    single-precision size estimate included, which follows the reference's 1 MiB compressed reads
    (csrc/workhorse.c:883-884) -- against the CPU checker; and identical with one zlib stream per file."""
    import gzip
    from kvarq_b200.engine import Engine
    g = synth.genome(300000, seed=81)
    ts = synth.mtbc_templates(g, seed=82, n_phylo=6, n_res=10, regions=(3, 81), n_spacers=5)
    iso = synth.isolate(g, ts, seed=83)
    seqs = ts.seqs()
    fq1 = synth.fastq_reads(iso, 110000, readlen=250, seed=84, qmean0=37.0, qdrop=20.0).tobytes()
    fq2 = synth.fastq_reads(iso, 70000, readlen=150, seed=85, mate=2).tobytes()
    f1 = str(tmp_path / 'big_1.fastq.gz')
    open(f1, 'wb').write(gzip.compress(fq1, 6))
    cut = [fq2.index(b'\n@SYN', int(len(fq2) * x)) + 1 for x in (0.3, 0.31)]
    f2 = str(tmp_path / 'big_2.fastq.gz')
    open(f2, 'wb').write(gzip.compress(fq2[:cut[0]], 6) + gzip.compress(fq2[cut[0]:cut[1]], 1) + gzip.compress(fq2[cut[1]:], 9) + b'tail')
    assert os.path.getsize(f1) > (12 << 20)
    e = Engine(0)
    e.config(nthreads=8, **DEFAULT)
    want = oracle(checker, [f1, f2], seqs, DEFAULT, impl='oracle', nthreads=os.cpu_count() or 8)
    assert want['stats']['records_parsed'] == 180000 and len(want['hits']) > 300
    got = e.findseqs([f1, f2], seqs, nplus=ts.nplus)
    assert_same_result(got, want, what='parallel inflate')
    monkeypatch.setenv('KVQ_GZ_THREADS', '1')
    one = e.findseqs([f1, f2], seqs, nplus=ts.nplus)
    assert_same_result(one, want, what='one zlib stream per file')
    if checker.available('ref'):
        small = str(tmp_path / 'ref_1.fastq.gz')
        open(small, 'wb').write(gzip.compress(fq1[:fq1.index(b'\n@SYN', 30 << 20) + 1], 6))
        monkeypatch.delenv('KVQ_GZ_THREADS')
        got = e.findseqs(small, seqs[:8])
        assert_same_result(got, oracle(checker, small, seqs[:8], DEFAULT, impl='ref'), what='parallel inflate vs the reference C')
    e.close()


def test_bgzf_members_inflated_side_by_side(checker, tmp_path, monkeypatch):
    """A .fastq.gz made of 64 KiB members (BGZF, what Illumina's converters write): the member rounds of
    gz_parallel.h decode the members side by side; hits, hitseqs and stats (incl. the size estimate, which moves
    with every member: csrc/workhorse.c:842-884) against the CPU checker and against one zlib stream."""
    import gzip
    from kvarq_b200.engine import Engine
    g = synth.genome(300000, seed=91)
    ts = synth.mtbc_templates(g, seed=92, n_phylo=6, n_res=10, regions=(3, 81), n_spacers=5)
    iso = synth.isolate(g, ts, seed=93)
    seqs = ts.seqs()
    fq = synth.fastq_reads(iso, 120000, readlen=150, seed=94).tobytes()                 # ~40 MB, ~600 members
    fn = str(tmp_path / 'blocks.fastq.gz')
    open(fn, 'wb').write(b''.join(gzip.compress(fq[i:i + 65280], 6) for i in range(0, len(fq), 65280)))
    e = Engine(0)
    e.config(nthreads=8, **DEFAULT)
    want = oracle(checker, fn, seqs, DEFAULT, impl='oracle', nthreads=os.cpu_count() or 8)
    assert want['stats']['records_parsed'] == 120000 and len(want['hits']) > 200
    got = e.findseqs(fn, seqs, nplus=ts.nplus)
    assert_same_result(got, want, what='member rounds')
    monkeypatch.setenv('KVQ_GZ_THREADS', '1')
    assert_same_result(e.findseqs(fn, seqs, nplus=ts.nplus), want, what='one zlib stream')
    e.close()
