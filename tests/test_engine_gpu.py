"""Parity tests proper: the CUDA engine, called through the C ABI (kvarq_b200.engine ->
libkvarq_b200.so), against the CPU checkers and the committed golden vectors.  Bit-exact:
hits (reference nthreads=1 order), hitseqs, stats, fused coverage.

The first half mirrors the reference's tests/test_engine.py test by test.
"""
import gzip
import os
import random
import threading

import numpy as np
import pytest

from util import assert_same_result, coverage_from_hits
from kvarq_b200 import synth
from kvarq_b200.fastq import FastqFileFormatException

pytestmark = pytest.mark.gpu

ENGINE_SEQS = ("CCC", "TTTT", "TATATATA", "TGTAG", "ATATT",
               "GAGCATGTGGAGCAACTTGTGGGAGCGCCGGGCAACGCCCTGTCTCTTAT",
               "...NACTTCCTCTCTACTGGTGTCGGCGGTGAAAGAGCTTACGTACTCTTCGAT...")
DEFAULT = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')


def oracle(ck, files, seqs, cfg, impl=None):
    impl = impl or ('ref' if ck.available('ref') else 'oracle')
    return ck.findseqs(files, seqs, nthreads=1, impl=impl, **cfg)


@pytest.fixture(autouse=True)
def _setup(gpu_engine):
    # EngineTest.setUp, tests/test_engine.py:133-135
    gpu_engine.config(nthreads=1, maxerrors=2, minoverlap=25, minreadlength=10, Amin='!', Azero='!')


# ----------------------------------------------------------------- tests/test_engine.py ports --
@pytest.mark.parametrize('gz', [False, True])
def test_findseqs(gpu_engine, ref_fastqs, gz):
    """test_findseqs / test_gz (tests/test_engine.py:137-181)"""
    engine = gpu_engine
    engine.config(maxerrors=0, minoverlap=1000, minreadlength=3, Amin='!')
    fname = os.path.join(ref_fastqs, 'test_engine.fastq' + ('.gz' if gz else ''))
    hits = engine.findseqs(fname, ENGINE_SEQS)['hits']
    data = gzip.open(fname, 'rb').read() if gz else open(fname, 'rb').read()
    x = [0] * len(ENGINE_SEQS)
    for hit in hits:
        x[hit.seq_nr] += 1
        seq = ENGINE_SEQS[hit.seq_nr]
        if hit.seq_pos < 0:
            bps = data[hit.file_pos - hit.seq_pos:hit.file_pos - hit.seq_pos + hit.length]
        else:
            bps = data[hit.file_pos:hit.file_pos + hit.length]
            seq = seq[hit.seq_pos:hit.seq_pos + hit[3]]
        assert bps.decode() == seq
    assert x == [19, 1, 0, 1, 1, 1, 1]


@pytest.mark.parametrize('gz', [False, True])
def test_paired(gpu_engine, ref_fastqs, gz):
    """test_paired (tests/test_engine.py:184-205): the whole dict, stats included"""
    engine = gpu_engine
    engine.config(maxerrors=0, minoverlap=1000, minreadlength=3, Amin='!')
    sfx = '.gz' if gz else ''
    ret = engine.findseqs(os.path.join(ref_fastqs, 'test_engine.fastq' + sfx), ENGINE_SEQS)
    ret_12 = engine.findseqs((os.path.join(ref_fastqs, 'test_engine_1.fastq' + sfx),
                              os.path.join(ref_fastqs, 'test_engine_2.fastq' + sfx)), ENGINE_SEQS)
    assert dict(ret) == dict(ret_12)


def test_maxerror(gpu_engine, ref_fastqs):
    engine = gpu_engine
    engine.config(minreadlength=25, minoverlap=25, Amin='!')
    seqs = ("CAGCATGTGGAGCAACTTGTGGGAGCGCCGGGCAACGCCCTGTCTCTTAT",
            "CTGCATGTGGAGCAACTTGTGGGAGCGCCGGGCAACGCCCTGTCTCTTAT",
            "CTCCATGTGGAGCAACTTGTGGGAGCGCCGGGCAACGCCCTGTCTCTTAT")
    for maxerrors in range(4):
        engine.config(maxerrors=maxerrors)
        hits = engine.findseqs(os.path.join(ref_fastqs, 'test_engine.fastq'), seqs)['hits']
        assert len(hits) == maxerrors


def test_minoverlap(gpu_engine, ref_fastqs):
    engine = gpu_engine
    fname = os.path.join(ref_fastqs, 'test_engine.fastq')
    seqs = ("TCGATGCGATCTGTCAAGTCGGTGGCGGTA...", "TCGATGCGATCTG.CAAGTCGGTGGCGGTA...",
            "...NTGAACGTATCGCCTCGAGGGACTT", "...NTGAACGTATCG.CTCGAGGGACTT")
    engine.config(maxerrors=0, minreadlength=25, minoverlap=30, Amin='!')
    hits = engine.findseqs(fname, seqs)['hits']
    assert len(hits) == 1 and hits[0].seq_nr == 0 and hits[0].seq_pos < 0
    engine.config(maxerrors=0, minoverlap=25)
    hits = engine.findseqs(fname, seqs)['hits']
    assert len(hits) == 2
    for hit in hits:
        assert hit[0] != 3 or hit[2] > 0
    engine.config(maxerrors=1, minoverlap=25)
    assert len(engine.findseqs(fname, seqs)['hits']) == 4


def test_Amin(gpu_engine, ref_fastqs):
    engine = gpu_engine
    fname = os.path.join(ref_fastqs, 'test_engine.fastq')
    engine.config(Amin='H', minreadlength=4, maxerrors=0)
    ret = engine.findseqs(fname, ("GGAG", "CCGAC"))
    assert len(ret['hits']) == 1
    assert ret['stats']['readlengths'][5] == 3
    assert ret['stats']['readlengths'][4] == 5
    engine.config(Amin='G')
    assert len(engine.findseqs(fname, ("GGAG", "CCGAC"))['hits']) == 2


def test_hits(gpu_engine, tmp_path):
    """test_hits (tests/test_engine.py:274-321): 100 reads of 10+100+10 bp covering a 51-mer"""
    engine = gpu_engine
    rng = random.Random(7)
    rs = lambda n: ''.join(rng.choice('ACGT') for _ in range(n))
    seq = rs(51)
    minoverlap, readlength, n = 25, 100, 100
    recs = []
    for i in range(n):
        overlap = rng.randint(minoverlap, readlength)
        if overlap > len(seq):
            l = rng.randint(0, readlength - len(seq))
            seqx = rs(l) + seq + rs(readlength - len(seq) - l)
        elif rng.random() < .5:
            seqx = rs(readlength - overlap) + seq[:overlap]
        else:
            seqx = seq[-overlap:] + rs(readlength - overlap)
        bases = rs(10) + seqx + rs(10)
        phred = ''.join(chr(33 + rng.randint(0, 12)) for _ in range(10)) \
            + ''.join(chr(33 + rng.randint(13, 40)) for _ in range(readlength)) \
            + ''.join(chr(33 + rng.randint(0, 12)) for _ in range(10))
        recs.append('@GENERATED%010d\n%s\n+\n%s\n' % (i, bases, phred))
    fn = str(tmp_path / 'hits.fastq')
    open(fn, 'w').write(''.join(recs))
    engine.config(nthreads=3, Amin=chr(33 + 13), maxerrors=0,
                  minreadlength=rng.randint(minoverlap, readlength), minoverlap=minoverlap)
    ret = engine.findseqs(fn, [seq])
    assert ret['stats']['readlengths'][readlength] == n
    assert len(ret['hits']) == n
    seqx = ''.join([i % minoverlap != 0 and b or {'A': 'C', 'C': 'G', 'G': 'T', 'T': 'A'}[b]
                    for i, b in enumerate(seq)])
    ret = engine.findseqs(fn, [seqx])
    assert ret['stats']['readlengths'][readlength] == n
    assert len(ret['hits']) == 0


def test_fastq(gpu_engine, tmp_path, checker):
    """test_fastq (tests/test_engine.py:324-346) + the exact message of the reference"""
    engine = gpu_engine
    fn = str(tmp_path / 'bad.fastq')
    for text in ('_IDENTIFIER\nACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGT\n+\n'
                 '#############################################\n',
                 '@IDENTIFIER\nACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGT\n-\n'
                 '#############################################\n'):
        open(fn, 'w').write(text)
        with pytest.raises(FastqFileFormatException) as e:
            engine.findseqs(fn, [])
        with pytest.raises(checker.OracleError) as o:
            oracle(checker, fn, [], DEFAULT)
        assert str(e.value) == o.value.msg


def test_forward_fastq(gpu_engine, tmp_path):
    """test_forward_fastq (tests/test_engine.py:349-359): '+IDENTIFIER' lines and CRLF"""
    engine = gpu_engine
    engine.config(Amin='#', nthreads=2, minoverlap=80)
    fn = str(tmp_path / 'fw.fastq')
    for n in [3, 5, 7, 133]:
        for plus in ['+', '+IDENTIFIER']:
            for cr in ['\n', '\r\n']:
                record = '@IDENTIFIER' + cr + 'A' * 80 + cr + plus + cr + '#' * 80 + cr
                open(fn, 'wb').write((record * n).encode())
                ret = engine.findseqs(fn, ['A' * 80])
                assert len(ret['hits']) == n


# ------------------------------------------------------------------------ golden vectors ------
def test_golden_cases(gpu_engine, ref_fastqs, golden):
    for case in golden['cases']:
        gpu_engine.config(nthreads=1, **case['cfg'])
        files = [os.path.join(ref_fastqs, f) for f in case['files']]
        got = gpu_engine.findseqs(files if len(files) > 1 else files[0], case['seqs'])
        assert_same_result(got, case, what=case['name'])


# --------------------------------------------------------------- differential vs the checker --
def _random_fastq(rng, path, nrec, maxlen, alpha, quals='!#5II'):
    recs = []
    for i in range(nrec):
        L = rng.randint(1, maxlen)
        s = ''.join(rng.choice(alpha) for _ in range(L))
        q = ''.join(rng.choice(quals) for _ in range(L))
        recs.append('@r%d\n%s\n+\n%s\n' % (i, s, q))
    open(path, 'w').write(''.join(recs))


def test_random_small_configs(gpu_engine, checker, tmp_path):
    """SURVEY Appendix B differential: tiny reads/templates, every rule incl. the duplicate quirk
    (exhaustive matcher on the GPU)."""
    rng = random.Random(1234)
    fn = str(tmp_path / 'r.fastq')
    dups = 0
    for trial in range(120):
        E = rng.choice([0, 1, 2, 3]); O = rng.choice([1, 3, 5, 8, 12, 1000]); R = rng.choice([1, 3, 6])
        alpha = rng.choice(['AC', 'ACGT', 'ACGTN'])
        _random_fastq(rng, fn, rng.randint(1, 40), 40, alpha)
        seqs = [''.join(rng.choice(alpha) for _ in range(rng.randint(1, 30))) for _ in range(rng.randint(1, 6))]
        cfg = dict(maxerrors=E, minoverlap=O, minreadlength=R, Amin='5')
        gpu_engine.config(**cfg)
        got = gpu_engine.findseqs(fn, seqs)
        want = oracle(checker, fn, seqs, cfg)
        assert_same_result(got, want, what='trial %d %r' % (trial, cfg))
        h = [tuple(x) for x in want['hits']]
        dups += len(h) - len(set(h))
    assert dups > 0, 'the flush-right duplicate quirk was never exercised'


def test_random_seeded_configs(gpu_engine, checker, tmp_path):
    """Reads and templates long enough for the seeded path (k=8), low-complexity alphabets so
    that partial overlaps, multiple seeds per diagonal and near-misses are frequent."""
    rng = random.Random(99)
    fn = str(tmp_path / 's.fastq')
    for trial in range(70 * int(os.environ.get('KVQ_STRESS', 1))):     # KVQ_STRESS=n: n times the trials
        E = rng.choice([0, 1, 2, 3]); O = rng.choice([8 * (E + 1), 25, 30, 40]); R = rng.choice([10, 25, 40])
        alpha = rng.choice(['AC', 'ACGT', 'ACGTN', 'ACGTacgt.'])
        master = ''.join(rng.choice(alpha) for _ in range(400))
        seqs = []
        for _ in range(rng.randint(1, 8)):
            a = rng.randint(0, 340); ln = rng.randint(8 * (E + 1), 140)
            s = list(master[a:a + ln])
            for _ in range(rng.randint(0, 3)):
                s[rng.randrange(len(s))] = rng.choice(alpha)
            seqs.append(''.join(s))
        recs = []
        for i in range(rng.randint(20, 120)):
            L = rng.randint(20, 160)
            a = rng.randint(0, 400 - 20)
            s = list((master[a:a + L] + ''.join(rng.choice(alpha) for _ in range(L)))[:L])
            for _ in range(rng.randint(0, 4)):
                s[rng.randrange(L)] = rng.choice(alpha)
            q = ''.join(rng.choice('5IIIIIII#') for _ in range(L))
            recs.append('@r%d\n%s\n+\n%s\n' % (i, ''.join(s), q))
        open(fn, 'w').write(''.join(recs))
        cfg = dict(maxerrors=E, minoverlap=O, minreadlength=R, Amin='5')
        gpu_engine.config(**cfg)
        got = gpu_engine.findseqs(fn, seqs)
        want = oracle(checker, fn, seqs, cfg)
        assert_same_result(got, want, what='trial %d %r' % (trial, cfg))


@pytest.fixture(scope='module')
def multi_chunk(tmp_path_factory):
    d = tmp_path_factory.mktemp('gpu_multichunk')
    g = synth.genome(200000, seed=5)
    ts = synth.mtbc_templates(g, seed=3, n_phylo=6, n_res=10, regions=(3, 81), n_spacers=5)
    fq = synth.fastq_reads(synth.isolate(g, ts), 11000, readlen=150, seed=9).tobytes()
    plain = os.path.join(str(d), 'mc.fastq')
    open(plain, 'wb').write(fq)
    gz1 = os.path.join(str(d), 'mc1.fastq.gz')
    with gzip.open(gz1, 'wb', compresslevel=6) as f:
        f.write(fq)
    gzm = os.path.join(str(d), 'mcm.fastq.gz')
    cut = fq.index(b'\n@SYN', len(fq) // 3) + 1
    open(gzm, 'wb').write(gzip.compress(fq[:cut], 6) + gzip.compress(fq[cut:], 6))
    return dict(plain=plain, gz1=gz1, gzm=gzm, seqs=ts.seqs(), nplus=ts.nplus, bytes=fq)


@pytest.mark.parametrize('kind', ['plain', 'gz1', 'gzm'])
def test_multi_chunk_streams(gpu_engine, checker, multi_chunk, kind):
    """4 reference chunks, plain / gzip / multi-member gzip: hits, stats (incl. the float size
    estimate of gz input) identical to the reference"""
    mc = multi_chunk
    gpu_engine.config(**DEFAULT)
    got = gpu_engine.findseqs(mc[kind], mc['seqs'])
    want = oracle(checker, mc[kind], mc['seqs'], DEFAULT)
    assert len(want['hits']) > 50
    assert_same_result(got, want, what=kind)


def test_many_gz_files_inflated_side_by_side(gpu_engine, checker, multi_chunk, tmp_path):
    """.gz files of one scan are inflated ahead by one host thread each (up to 4 at a time): five
    files -- single-member, multi-member, one tiny, one empty stream -- must give the reference's
    hits and stats (continuing file_pos, float size estimate from the reference's file position)."""
    fq = multi_chunk['bytes']
    cuts = [0]
    for frac in (0.17, 0.45, 0.46, 0.8):
        cuts.append(fq.index(b'\n@SYN', int(len(fq) * frac)) + 1)
    cuts.append(len(fq))
    files = []
    for k in range(5):
        part = fq[cuts[k]:cuts[k + 1]]
        fn = str(tmp_path / ('lane%d.fastq.gz' % k))
        if k == 1:
            half = part.index(b'\n@SYN', len(part) // 2) + 1
            open(fn, 'wb').write(gzip.compress(part[:half], 6) + gzip.compress(b'', 6) + gzip.compress(part[half:], 1))
        else:
            open(fn, 'wb').write(gzip.compress(part, 6))
        files.append(fn)
    gpu_engine.config(**DEFAULT)
    got = gpu_engine.findseqs(files, multi_chunk['seqs'])
    want = oracle(checker, files, multi_chunk['seqs'], DEFAULT)
    assert len(want['hits']) > 50
    assert_same_result(got, want, what='five gz files')
    open(files[2], 'wb').write(b'not a gzip file')
    with pytest.raises(IOError):
        gpu_engine.findseqs(files, multi_chunk['seqs'])


def test_fused_coverage(gpu_engine, checker, multi_chunk):
    """K6: coverage + alternate-base counts == Coverage.apply_hit over the reference's hits"""
    from kvarq_b200.engine import Engine
    mc = multi_chunk
    e = Engine(0)
    e.config(**DEFAULT)
    got = e.findseqs(mc['plain'], mc['seqs'], nplus=mc['nplus'])
    want = oracle(checker, mc['plain'], mc['seqs'], DEFAULT)
    cov, mut = coverage_from_hits(want['hits'], want['hitseqs'], mc['seqs'], mc['nplus'])
    assert got.coverage.tolist() == cov
    assert got.mutcount.tolist() == mut
    assert sum(cov) == sum(want['stats']['nseqbasehits'])
    e.close()


def test_mem_and_staged_paths(gpu_engine, checker, multi_chunk):
    from kvarq_b200.engine import Engine
    mc = multi_chunk
    e = Engine(0)
    e.config(**DEFAULT)
    want = oracle(checker, mc['plain'], mc['seqs'], DEFAULT)
    a = e.findseqs_mem([mc['bytes']], mc['seqs'])
    assert_same_result(a, want, what='mem')
    half = mc['bytes'].index(b'\n@SYN', len(mc['bytes']) // 2) + 1
    b = e.findseqs_mem([mc['bytes'][:half], mc['bytes'][half:]], mc['seqs'])
    assert [tuple(h) for h in b['hits']] == [tuple(h) for h in want['hits']]
    st = e.stage([np.frombuffer(mc['bytes'], dtype=np.uint8)])
    c = e.scan_staged(st, mc['seqs'])
    assert_same_result(c, want, what='staged')
    d = e.scan_staged(st, mc['seqs'], want_hits=False)
    assert d['stats']['nseqhits'] == want['stats']['nseqhits'] and d.timing['kernel_ms'] > 0
    assert d.timing['bases'] == synth.count_bases(mc['bytes'])
    st.free(); e.close()


def test_long_and_odd_records(gpu_engine, checker, tmp_path):
    """records beyond a tile's spill margin, reads beyond the packed slot, empty lines,
    quality line longer than the read, tiny and empty files"""
    rng = random.Random(5)
    rs = lambda n: ''.join(rng.choice('ACGT') for _ in range(n))
    tpl = rs(60)
    recs = []
    for i in range(300):
        L = rng.choice([30, 100, 600, 1100, 3000, 12000]) if i % 7 == 0 else rng.randint(30, 200)
        s = list(rs(L)); a = rng.randint(0, L - 25); s[a:a + 60] = list(tpl); s = ''.join(s)[:L]
        recs.append('@r%d %s\n%s\n+\n%s\n' % (i, 'x' * rng.randint(0, 80), s, 'I' * L))
    recs.insert(5, '@e\n\n+\n\n')
    recs.insert(9, '@q\nACGTACGTACGTACGTACGTACGTACGTAC\n+\n' + 'I' * 45 + '\n')
    fn = str(tmp_path / 'long.fastq')
    open(fn, 'w').write(''.join(recs))
    cfg = dict(maxerrors=1, minoverlap=20, minreadlength=20, Amin='5')
    gpu_engine.config(**cfg)
    got = gpu_engine.findseqs(fn, [tpl, synth.revcomp(tpl.encode()).decode()])
    want = oracle(checker, fn, [tpl, synth.revcomp(tpl.encode()).decode()], cfg)
    assert len(want['hits']) > 100
    assert_same_result(got, want, what='long records')
    for text in ('', '@only\nACGT\n', '@a\nACGTACGTAC\n+\nIIIIIIIIII\n@b\nACGTACGTAC\n+\nIIIIIIIIII'):
        open(fn, 'w').write(text)
        gpu_engine.config(maxerrors=0, minoverlap=1000, minreadlength=5, Amin='!')
        got = gpu_engine.findseqs(fn, ['ACGTACGTAC'])
        want = oracle(checker, fn, ['ACGTACGTAC'], dict(maxerrors=0, minoverlap=1000, minreadlength=5, Amin='!'))
        assert_same_result(got, want, what='tiny %r' % text[:10])


def test_line_phase_inference(gpu_engine, checker, tmp_path, monkeypatch):
    """A tile takes its line phase (newlines of the segment before it, mod 4) from its own line
    starts when only one phase makes every line 0 start with '@' and every line 2 with '+', and
    checks it against the newline counts afterwards.  Well-formed streams whose quality lines
    start with '@' / '+' (inference not decisive: the tile waits for the counts), and malformed
    ones (a missing or an extra line shifts the phase of everything behind it: the batch is run
    again with exact phases) must give the reference's result or the reference's exception."""
    rng = random.Random(11)
    rs = lambda n: ''.join(rng.choice('ACGT') for _ in range(n))
    tpl = rs(60)
    seqs = [tpl, synth.revcomp(tpl.encode()).decode()]

    def records(n, q0=lambda i: None):
        out = []
        for i in range(n):
            L = rng.randint(60, 160)
            sq = list(rs(L))
            if i % 5 == 0:
                a = rng.randint(0, L - 30); sq[a:a + 60] = list(tpl)
            sq = ''.join(sq)[:L]
            q = ''.join(rng.choice('5?II') for _ in range(L))
            if q0(i):
                q = q0(i) + q[1:]
            out.append('@r%d\n%s\n+\n%s\n' % (i, sq, q))
        return out

    cases = {}
    cases['clean'] = records(3000)
    cases['quality starts with @'] = records(3000, lambda i: '@')
    cases['quality starts with +'] = records(3000, lambda i: '+')
    cases['some quality lines start with @ or +'] = records(3000, lambda i: '@+'[i % 2] if i % 3 == 0 else None)
    r = records(3000); r[1700] = r[1700].replace('\n+\n', '\n'); cases['a record without its + line'] = r
    r = records(3000); r[900] = r[900] + '\n'; cases['an empty line between two records'] = r
    r = records(3000); r[2500] = '_' + r[2500][1:]; cases['a late record without @'] = r
    r = records(3000); r.append('garbage\n+\nmore garbage'); cases['junk partial record at the end'] = r
    r = records(3000, lambda i: '@'); r[1200] = r[1200].replace('\n+\n', '\n-\n'); cases['bad + line among @ quality lines'] = r
    cfg = dict(maxerrors=1, minoverlap=20, minreadlength=20, Amin='5')
    fn = str(tmp_path / 'phase.fastq')
    for name, recs in cases.items():
        open(fn, 'w').write(''.join(recs))
        try:
            want, werr = oracle(checker, fn, seqs, cfg), None
        except checker.OracleError as o:
            want, werr = None, o.msg
        for exact in ('0', '1'):
            monkeypatch.setenv('KVQ_EXACT_PHASE', exact)
            gpu_engine.config(**cfg)
            what = '%s (exact phases: %s)' % (name, exact)
            if werr is None:
                got = gpu_engine.findseqs(fn, seqs)
                assert len(want['hits']) > 300, what
                assert_same_result(got, want, what=what)
            else:
                with pytest.raises(FastqFileFormatException) as e:
                    gpu_engine.findseqs(fn, seqs)
                assert str(e.value) == werr, what
    monkeypatch.delenv('KVQ_EXACT_PHASE')


def test_many_hits_overflow_buffers(gpu_engine, checker, tmp_path):
    """poly-A reads against a poly-A template: far more hits than reads; exercises the hit
    buffer regrow + re-run path"""
    fn = str(tmp_path / 'polya.fastq')
    open(fn, 'w').write(''.join('@p%d\n%s\n+\n%s\n' % (i, 'A' * 150, 'I' * 150) for i in range(4000)))
    cfg = dict(maxerrors=0, minoverlap=25, minreadlength=25, Amin='!')
    gpu_engine.config(**cfg)
    got = gpu_engine.findseqs(fn, ['A' * 60])
    want = oracle(checker, fn, ['A' * 60], cfg, impl='oracle')
    assert len(want['hits']) > (1 << 18)
    assert_same_result(got, want, what='poly-A')


def test_staged_hits_ordered_on_device(checker, tmp_path):
    """HBM-resident scans put their hits into the reference's order on the device (bucket sort over
    the stream position, ties -- several hits of one read, duplicate hits -- by template / rule /
    shift): repeated scans (result copies sized by the previous scan), a template set that makes
    most reads hit several templates, more hits than the device path takes (host sort), no hits."""
    from kvarq_b200.engine import Engine
    rng = random.Random(23)
    rs = lambda n: ''.join(rng.choice('ACGT') for _ in range(n))
    core = rs(90)
    tpls = [core[a:a + 45] for a in (0, 10, 20, 30, 45)] + [core]
    seqs = tpls + [synth.revcomp(t.encode()).decode() for t in tpls]
    recs = []
    for i in range(6000):
        L = rng.randint(40, 150)
        sq = list(rs(L))
        if i % 2 == 0:
            a = rng.randint(0, max(0, L - 40)); b0 = rng.randint(0, 50)
            sq[a:a + 90 - b0] = list(core[b0:]); sq = sq[:L]
        recs.append('@r%d\n%s\n+\n%s\n' % (i, ''.join(sq), 'I' * len(sq)))
    cfg = dict(maxerrors=1, minoverlap=20, minreadlength=20, Amin='5')
    e = Engine(0)
    e.config(**cfg)
    cases = {'several templates per read': (''.join(recs), seqs),
             'no hits': (''.join(recs[1::2]), [rs(60)]),
             'more hits than the device orders': (''.join('@p%d\n%s\n+\n%s\n' % (i, 'A' * 150, 'I' * 150) for i in range(1500)), ['A' * 60])}
    for name, (text, sq) in cases.items():
        fn = str(tmp_path / 'ord.fastq')
        open(fn, 'w').write(text)
        want = oracle(checker, fn, sq, cfg, impl='oracle' if 'more hits' in name else None)
        if 'several' in name:
            assert len(want['hits']) > 5000
        if 'more hits' in name:
            assert len(want['hits']) > 65536
        st = e.stage([np.frombuffer(text.encode(), dtype=np.uint8)])
        for rep in range(3):
            got = e.scan_staged(st, sq, nplus=0)
            assert_same_result(got, want, what='%s, scan %d' % (name, rep))
        st.free()
    # few and many hits in turn on the same stream (buffers and tables of the ordering kernels are reused from scan to scan)
    text = ''.join(recs[:2000]) + ''.join('@p%d\n%s\n+\n%s\n' % (i, 'A' * 150, 'I' * 150) for i in range(200))
    fn = str(tmp_path / 'ord2.fastq')
    open(fn, 'w').write(text)
    st = e.stage([np.frombuffer(text.encode(), dtype=np.uint8)])
    want_few = oracle(checker, fn, seqs, cfg)
    want_many = oracle(checker, fn, ['A' * 60], cfg, impl='oracle')
    assert 1000 < len(want_few['hits']) < 24576 < len(want_many['hits']) < 65536
    for rep, (sq, want) in enumerate([(seqs, want_few), (seqs, want_few), (['A' * 60], want_many), (['A' * 60], want_many), (seqs, want_few)]):
        assert_same_result(e.scan_staged(st, sq, nplus=0), want, what='few / many hits in turn, scan %d' % rep)
    st.free()
    e.close()


def test_survivor_queue_overflow(checker, multi_chunk, tmp_path, monkeypatch):
    """The scan kernel hands candidates to the emit kernel through a queue; when it overflows
    the emit / long-record kernels do nothing and the batch is run again with a queue that fits
    (file path and staged path, fused coverage must not be counted twice)."""
    from kvarq_b200.engine import Engine
    monkeypatch.setenv('KVQ_SURV_CAP', '16')
    e = Engine(0)
    e.config(**DEFAULT)
    fn, seqs = multi_chunk['plain'], multi_chunk['seqs']
    nplus = len(seqs) // 2
    want = oracle(checker, fn, seqs, DEFAULT)
    got = e.findseqs(fn, seqs, nplus=nplus)
    assert_same_result(got, want, what='survivor overflow, file')
    cov, mut = coverage_from_hits(want['hits'], want['hitseqs'], seqs, nplus)
    assert got.coverage.tolist() == cov and got.mutcount.tolist() == mut
    st = e.stage([np.frombuffer(open(fn, 'rb').read(), dtype=np.uint8)])
    for _ in range(2):
        got = e.scan_staged(st, seqs, nplus=nplus)
        assert_same_result(got, want, what='survivor overflow, staged')
        assert got.coverage.tolist() == cov and got.mutcount.tolist() == mut
    st.free()
    e.close()


def test_api_errors_and_state(gpu_engine, tmp_path):
    engine = gpu_engine
    engine.config(maxerrors=1, minoverlap=33, minreadlength=44, nthreads=5, Amin='+', Azero='@')
    assert engine.get_config() == {'maxerrors': 1, 'minoverlap': 33, 'minreadlength': 44,
                                   'nthreads': 5, 'Amin': '+', 'Azero': '@'}
    with pytest.raises(IOError):
        engine.findseqs(str(tmp_path / 'does_not_exist.fastq'), ['ACGT'])
    bad = str(tmp_path / 'bad.fastq.gz')
    open(bad, 'wb').write(b'this is not gzip')
    with pytest.raises(IOError):
        engine.findseqs(bad, ['ACGT'])
    with pytest.raises(TypeError):
        engine.findseqs(12, ['ACGT'])
    with pytest.raises(TypeError):
        engine.findseqs(bad, 12)
    with pytest.raises(TypeError):
        engine.config(bogus=1)


def test_stats_and_stop_from_another_thread(gpu_engine, tmp_path):
    """engine.stats() polled and engine.stop() called while findseqs runs (kvarq/cli.py:134-175)"""
    engine = gpu_engine
    g = synth.genome(100000, seed=8)
    fq = synth.fastq_reads(g, 300000, readlen=150, seed=3).tobytes()      # ~96 MB: two batches
    fn = str(tmp_path / 'big.fastq')
    open(fn, 'wb').write(fq)
    engine.config(**DEFAULT)
    out = {}
    t = threading.Thread(target=lambda: out.update(r=engine.findseqs(fn, [g[1000:1051].tobytes()])))
    t.start()
    seen = []
    while t.is_alive():
        st = engine.stats()
        seen.append(st['parsed'])
        assert set(st) == {'readlengths', 'progress', 'nseqbasehits', 'nseqhits', 'parsed', 'total',
                           'sigints', 'records_parsed'}
        assert 0.0 <= st['progress'] <= 1.0
        t.join(0.01)
    full = out['r']
    assert full['stats']['records_parsed'] == 300000 and full['stats']['parsed'] == len(fq)
    assert seen == sorted(seen)
    # stop: partial result, no exception
    t = threading.Thread(target=lambda: out.update(s=engine.findseqs(fn, [g[1000:1051].tobytes()])))
    t.start()
    engine.stop()
    t.join()
    assert out['s']['stats']['records_parsed'] <= 300000


def test_thousands_of_templates(gpu_engine, checker, tmp_path):
    """BASELINE config 4 at test scale: a few thousand 51-bp templates (packed template codes no
    longer fit shared memory, seed buckets overflow into the entry array), reads as in config 2."""
    g = synth.genome(120000, seed=41)
    rng = np.random.RandomState(42)
    comp = bytes.maketrans(b'ACGT', b'TGCA')
    plus = []
    for p in sorted(rng.choice(len(g) - 60, 1500, replace=False)):
        plus.append(bytes(g[p:p + 51]))
    seqs = [s.decode() for s in plus] + [s.translate(comp)[::-1].decode() for s in plus]
    fn = str(tmp_path / 'c4.fastq')
    open(fn, 'wb').write(synth.fastq_reads(g, 1500, readlen=150, seed=43).tobytes())
    from kvarq_b200.engine import Engine
    e = Engine(0)
    e.config(**DEFAULT)
    got = e.findseqs(fn, seqs, nplus=len(plus))
    want = oracle(checker, fn, seqs, DEFAULT, impl='oracle')
    assert_same_result(got, want, what='3000 templates')
    assert len(want['hits']) > 500
    assert got.timing['exhaustive_reads'] == 0
    cov, mut = coverage_from_hits(want['hits'], want['hitseqs'], seqs, len(plus))
    assert got.coverage.tolist() == cov and got.mutcount.tolist() == mut
    e.close()


def test_huge_seed_buckets(gpu_engine, checker, tmp_path):
    """Hundreds of templates sharing their k-mers: buckets beyond the 8-bit count of the first-entry
    table (the count then sits in the overflow array), every read matching many templates."""
    rng = random.Random(7)
    core = ''.join(rng.choice('ACGT') for _ in range(60))
    seqs = []
    for i in range(300):
        s = list(core[:51])
        s[rng.randrange(51)] = rng.choice('ACGT')          # near-identical templates
        seqs.append(''.join(s))
    recs = []
    for i in range(40):
        L = rng.randint(60, 120)
        left = ''.join(rng.choice('ACGT') for _ in range(rng.randint(0, L - 55)))
        s = (left + core + ''.join(rng.choice('ACGT') for _ in range(L)))[:L]
        recs.append('@r%d\n%s\n+\n%s\n' % (i, s, 'I' * L))
    fn = str(tmp_path / 'hb.fastq')
    open(fn, 'w').write(''.join(recs))
    gpu_engine.config(**DEFAULT)
    got = gpu_engine.findseqs(fn, seqs)
    want = oracle(checker, fn, seqs, DEFAULT, impl='oracle')
    assert_same_result(got, want, what='huge buckets')
    assert len(want['hits']) > 3000


def test_large_set_mixed_template_classes(gpu_engine, checker, tmp_path):
    """The large-set kernel variant (more than 2048 T1 pieces: 12-mer second level, all bucket
    entries dealt to lanes) with every template class at once: 25-bp spacers (pieces shorter
    than 12 bases: no second level for their keys), 51-bp SNP templates, a few long regions;
    noisy reads (errors inside the 12-mers, heavy trimming: seeds closer than 12 bases to the
    read end) and two values of maxerrors."""
    g = synth.genome(40000, seed=51)
    rng = np.random.RandomState(52)
    comp = bytes.maketrans(b'ACGT', b'TGCA')
    plus = [bytes(g[p:p + 25]) for p in sorted(rng.choice(len(g) - 700, 500, replace=False))]
    plus += [bytes(g[p:p + 51]) for p in sorted(rng.choice(len(g) - 700, 500, replace=False))]
    plus += [bytes(g[p:p + n]) for p, n in ((1000, 131), (9000, 611), (20000, 53), (30000, 36))]
    seqs = [s.decode() for s in plus] + [s.translate(comp)[::-1].decode() for s in plus]
    fn = str(tmp_path / 'mixed.fastq')
    fq = [synth.fastq_reads(g, 1200, readlen=150, seed=53, qmean0=30.0, qdrop=26.0, qsd=5.0).tobytes(),
          synth.fastq_reads(g, 600, readlen=60, seed=54, qmean0=26.0, qdrop=14.0, qsd=6.0, first_index=5000).tobytes()]
    open(fn, 'wb').write(b''.join(fq))
    from kvarq_b200.engine import Engine
    e = Engine(0)
    for cfg in (DEFAULT, dict(DEFAULT, maxerrors=1, minoverlap=20)):
        e.config(**cfg)
        got = e.findseqs(fn, seqs, nplus=len(plus))
        want = oracle(checker, fn, seqs, cfg, impl='oracle')
        assert_same_result(got, want, what='mixed classes %r' % (cfg,))
        assert len(want['hits']) > 1000
        assert got.timing['exhaustive_reads'] == 0
        cov, mut = coverage_from_hits(want['hits'], want['hitseqs'], seqs, len(plus))
        assert got.coverage.tolist() == cov and got.mutcount.tolist() == mut
    e.close()


def test_large_set_huge_buckets(gpu_engine, checker, tmp_path):
    """Large-set variant with buckets of hundreds of entries (the count escape of the bucket head,
    one bucket spread over many rounds of 32 lanes)."""
    rng = random.Random(17)
    core = ''.join(rng.choice('ACGT') for _ in range(70))
    seqs = []
    for i in range(800):
        s = list(core[:51])
        s[rng.randrange(51)] = rng.choice('ACGT')
        if i % 3 == 0:
            s[rng.randrange(51)] = rng.choice('ACGT')
        seqs.append(''.join(s))
    recs = []
    for i in range(30):
        L = rng.randint(56, 130)
        left = ''.join(rng.choice('ACGT') for _ in range(rng.randint(0, L - 55)))
        s = list((left + core + ''.join(rng.choice('ACGT') for _ in range(L)))[:L])
        if i % 2:
            s[rng.randrange(L)] = rng.choice('ACGT')
        recs.append('@r%d\n%s\n+\n%s\n' % (i, ''.join(s), 'I' * L))
    fn = str(tmp_path / 'hb2.fastq')
    open(fn, 'w').write(''.join(recs))
    gpu_engine.config(**DEFAULT)
    got = gpu_engine.findseqs(fn, seqs)
    want = oracle(checker, fn, seqs, DEFAULT, impl='oracle')
    assert_same_result(got, want, what='large set, huge buckets')
    assert len(want['hits']) > 5000


def test_paired_end_250_gz(gpu_engine, checker, tmp_path):
    """BASELINE config 3 at test scale: 2 x 250 bp, MiSeq-like quality decay (real trimming), two
    gz files scanned as one stream with a continuing file_pos."""
    g = synth.genome(150000, seed=51)
    ts = synth.mtbc_templates(g, seed=52)
    iso = synth.isolate(g, ts, seed=53)
    files = []
    for k in (1, 2):
        fq = synth.fastq_reads(iso, 4000, readlen=250, seed=60 + k, qmean0=37.0 - 3 * (k - 1), qdrop=26.0, mate=k).tobytes()
        fn = str(tmp_path / ('pe_%d.fastq.gz' % k))
        with gzip.open(fn, 'wb', compresslevel=6) as f:
            f.write(fq)
        files.append(fn)
    gpu_engine.config(**DEFAULT)
    got = gpu_engine.findseqs(files, ts.seqs())
    want = oracle(checker, files, ts.seqs(), DEFAULT)
    assert_same_result(got, want, what='paired-end gz')
    assert len(set(want['stats']['readlengths'])) > 3, 'quality trimming was not exercised'
