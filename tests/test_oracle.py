"""Pins the CPU checkers: the plain-C restatement (oracle/kvarq_oracle.c) against the golden
vectors produced by the reference itself (tests/golden/make_golden.py), against the known
answers of the reference's tests/test_engine.py, and -- when oracle/_ref is present --
differentially against the reference C compiled unmodified."""
import gzip
import os
import random

import numpy as np
import pytest

from util import assert_same_result
from kvarq_b200 import synth


def run(ck, impl, files, seqs, cfg, nthreads=1):
    return ck.findseqs(files, seqs, nthreads=nthreads, impl=impl, **cfg)


def test_oracle_matches_golden(checker, ref_fastqs, golden):
    for case in golden['cases']:
        files = [os.path.join(ref_fastqs, f) for f in case['files']]
        got = run(checker, 'oracle', files, case['seqs'], case['cfg'])
        assert_same_result(got, case, what=case['name'])


def test_ref_build_matches_golden(checker, ref_fastqs, golden):
    if not checker.available('ref'):
        pytest.skip('oracle/_ref not built (no /root/reference here)')
    for case in golden['cases']:
        files = [os.path.join(ref_fastqs, f) for f in case['files']]
        got = run(checker, 'ref', files, case['seqs'], case['cfg'])
        assert_same_result(got, case, what=case['name'])


def test_known_answers_of_reference_tests(golden):
    """tests/test_engine.py:137-271 and tests/test_analyser.py:57-64, as counts."""
    by = {c['name']: c for c in golden['cases']}
    x = [0] * 7
    for h in by['findseqs']['hits']:
        x[h[0]] += 1
    assert x == [19, 1, 0, 1, 1, 1, 1]
    assert by['findseqs_gz']['hits'] == by['findseqs']['hits']
    for k in ('hits', 'hitseqs', 'stats'):
        assert by['paired'][k] == by['findseqs'][k]                 # test_paired: whole dict equal
    for e in range(4):
        assert len(by['maxerror_%d' % e]['hits']) == e
    h = by['minoverlap_30']['hits']
    assert len(h) == 1 and h[0][0] == 0 and h[0][2] < 0
    assert len(by['minoverlap_25']['hits']) == 2
    assert all(h[0] != 3 or h[2] > 0 for h in by['minoverlap_25']['hits'])
    assert len(by['minoverlap_25_e1']['hits']) == 4
    assert len(by['Amin_H']['hits']) == 1
    assert by['Amin_H']['stats']['readlengths'][5] == 3 and by['Amin_H']['stats']['readlengths'][4] == 5
    assert len(by['Amin_G']['hits']) == 2
    # spoligo octal 400000000000001 <=> only spacers 0 and 42 are seen (>= 2 hits each strand-summed)
    n = [0] * 43
    for h in by['analyser_spoligo']['hits']:
        n[h[0] % 43] += 1
    assert [i for i, c in enumerate(n) if c >= 2] == [0, 42]


def _random_case(rng, d, trial):
    E = rng.choice([0, 1, 2, 3]); O = rng.choice([1, 3, 5, 8, 12, 1000]); R = rng.choice([1, 3, 6])
    alpha = rng.choice(['AC', 'ACGT', 'ACGTN'])
    recs = []
    for i in range(rng.randint(1, 40)):
        L = rng.randint(1, 40)
        s = ''.join(rng.choice(alpha) for _ in range(L))
        q = ''.join(rng.choice('!#5II') for _ in range(L))
        recs.append('@r%d\n%s\n+\n%s\n' % (i, s, q))
    fn = os.path.join(d, 'rand%d.fastq' % trial)
    with open(fn, 'w') as f:
        f.write(''.join(recs))
    seqs = [''.join(rng.choice(alpha) for _ in range(rng.randint(1, 30))) for _ in range(rng.randint(1, 6))]
    return fn, seqs, dict(maxerrors=E, minoverlap=O, minreadlength=R, Amin='5')


def test_restatement_vs_reference_random(checker, tmp_path):
    if not checker.available('ref'):
        pytest.skip('oracle/_ref not built')
    rng = random.Random(20141009)
    for trial in range(150):
        fn, seqs, cfg = _random_case(rng, str(tmp_path), trial)
        a = run(checker, 'ref', fn, seqs, cfg)
        b = run(checker, 'oracle', fn, seqs, cfg, nthreads=3)
        assert_same_result(b, a, what='trial %d %r' % (trial, cfg))
        os.remove(fn)


@pytest.fixture(scope='module')
def multi_chunk(tmp_path_factory):
    """~3.4 MiB of synthetic reads (4 reference chunks) + gz single- and multi-member copies."""
    d = tmp_path_factory.mktemp('multichunk')
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
    with open(gzm, 'wb') as f:
        f.write(gzip.compress(fq[:cut], 6) + gzip.compress(fq[cut:], 6))
    return dict(plain=plain, gz1=gz1, gzm=gzm, seqs=ts.seqs(), nplus=ts.nplus, bytes=fq)


DEFAULT = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')


def test_multi_chunk_and_gz(checker, multi_chunk):
    mc = multi_chunk
    o = run(checker, 'oracle', mc['plain'], mc['seqs'], DEFAULT, nthreads=4)
    assert o['stats']['records_parsed'] == 11000 and len(o['hits']) > 50
    assert o['stats']['parsed'] == len(mc['bytes'])
    for k in ('gz1', 'gzm'):
        z = run(checker, 'oracle', mc[k], mc['seqs'], DEFAULT, nthreads=2)
        assert z['hits'] == o['hits'] and z['hitseqs'] == o['hitseqs']
        assert z['stats']['readlengths'] == o['stats']['readlengths']
    if checker.available('ref'):
        for k in ('plain', 'gz1', 'gzm'):
            r = run(checker, 'ref', mc[k], mc['seqs'], DEFAULT)
            x = run(checker, 'oracle', mc[k], mc['seqs'], DEFAULT)
            assert_same_result(x, r, what=k)


def test_trailing_record_without_newline_is_dropped(checker, tmp_path):
    fn = str(tmp_path / 't.fastq')
    open(fn, 'w').write('@a\nACGTACGTAC\n+\nIIIIIIIIII\n@b\nACGTACGTAC\n+\nIIIIIIIIII')
    r = run(checker, 'oracle', fn, ['ACGTACGTAC'], dict(maxerrors=0, minoverlap=1000, minreadlength=5, Amin='!'))
    assert r['stats']['records_parsed'] == 1 and len(r['hits']) == 1


def test_format_errors(checker, tmp_path):
    fn = str(tmp_path / 'bad.fastq')
    open(fn, 'w').write('_IDENTIFIER\nACGT\n+\n####\n')
    with pytest.raises(checker.OracleError) as e:
        run(checker, 'oracle', fn, [], DEFAULT)
    assert e.value.kind == 'FastqFileFormatException' and 'fpos=0' in e.value.msg
    open(fn, 'w').write('@IDENTIFIER\nACGT\n-\n####\n')
    with pytest.raises(checker.OracleError) as e:
        run(checker, 'oracle', fn, [], DEFAULT)
    assert e.value.kind == 'FastqFileFormatException' and 'fpos=17' in e.value.msg


def test_synth_is_deterministic():
    g = synth.genome(50000, seed=1)
    a = synth.fastq_reads(g, 1234, readlen=100, seed=2)
    b = synth.fastq_reads(g, 1234, readlen=100, seed=2)
    assert np.array_equal(a, b)
    assert synth.count_bases(a) == 123400
    ts = synth.mtbc_templates(synth.genome(300000, seed=2))
    assert ts.nplus == 132 and sum(len(s) for s in ts.plus) == 6318      # 6 318 bp per strand


def test_restatement_vs_reference_large_mixed_set(checker, tmp_path):
    """The large-set template shapes of the GPU tests (>= 2048 seed pieces: 25 / 51 / 611-bp classes
    mixed) are checked on the GPU against the restatement; here the restatement itself is pinned
    to the reference's own C on that shape."""
    if not checker.available('ref'):
        pytest.skip('oracle/_ref not built')
    g = synth.genome(40000, seed=51)
    rng = np.random.RandomState(52)
    comp = bytes.maketrans(b'ACGT', b'TGCA')
    plus = [bytes(g[p:p + 25]) for p in sorted(rng.choice(len(g) - 700, 250, replace=False))]
    plus += [bytes(g[p:p + 51]) for p in sorted(rng.choice(len(g) - 700, 450, replace=False))]
    plus += [bytes(g[p:p + n]) for p, n in ((1000, 131), (9000, 611), (20000, 53), (30000, 36))]
    seqs = [s.decode() for s in plus] + [s.translate(comp)[::-1].decode() for s in plus]
    assert 3 * len(seqs) >= 2048
    fn = str(tmp_path / 'mixed.fastq')
    fq = [synth.fastq_reads(g, 500, readlen=150, seed=53, qmean0=30.0, qdrop=26.0, qsd=5.0).tobytes(),
          synth.fastq_reads(g, 300, readlen=60, seed=54, qmean0=26.0, qdrop=14.0, qsd=6.0, first_index=5000).tobytes()]
    open(fn, 'wb').write(b''.join(fq))
    for cfg in (DEFAULT, dict(DEFAULT, maxerrors=1, minoverlap=20)):
        a = run(checker, 'ref', fn, seqs, cfg)
        b = run(checker, 'oracle', fn, seqs, cfg, nthreads=4)
        assert len(a['hits']) > 300
        assert_same_result(b, a, what='large mixed set %r' % (cfg,))


def test_restatement_vs_reference_negative_amin(checker, tmp_path):
    """Amin >= 0x80 (a negative signed char): the restatement follows the reference's signed
    compare (workhorse.c:1058), incl. runs left open by a '\\n' that is itself >= Amin."""
    if not checker.available('ref'):
        pytest.skip('oracle/_ref not built')
    rng = random.Random(3)
    g = synth.genome(20000, seed=5)
    tpl = bytes(g[3000:3051])
    recs = []
    for i in range(400):
        L = rng.randint(30, 120)
        a = rng.randrange(2900, 3100)
        q = bytes(rng.choices([0x85, 0x9f, 0xa0, 0xc8, 0xff, 0x49, 0x21], [1, 1, 5, 5, 5, 20, 3], k=L))
        recs.append(b'@n%d\n%s\n+\n%s\n' % (i, bytes(g[a:a + L]), q))
    fn = str(tmp_path / 'neg.fastq')
    open(fn, 'wb').write(b''.join(recs))
    for amin in ('\x80', '\xa0', '\xff'):
        cfg = dict(DEFAULT, Amin=amin)
        a = run(checker, 'ref', fn, [tpl], cfg)
        b = run(checker, 'oracle', fn, [tpl], cfg, nthreads=2)
        assert_same_result(b, a, what='Amin=%r' % amin)
    assert len(a['stats']['readlengths']) > 1
