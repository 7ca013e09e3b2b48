"""Consumer of the hit list: Coverage arithmetic as pinned by the reference's
tests/test_analyser.py:71-107, and calls from the fused kernel arrays == calls from the hits."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from kvarq_b200 import analyse, synth
from kvarq_b200.analyse import Coverage, Hit, Sequence

DEFAULT = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')


def test_coverage_reference_case():
    """tests/test_analyser.py:71-107, verbatim expectations"""
    #   AACCGGTT    : template
    #   ATCCGGTTTT  : hit1
    # AAAACCGGTT    : hit2
    #  AATCCGGTTA   : hit3
    cov = Coverage(Sequence('AACCGGTT'))
    cov.apply_hit(Hit(seq_nr=0, file_pos=-1, seq_pos=0, length=8, readlength=10), 'ATCCGGTTTT', on_plus_strand=True)
    assert cov.minf() == 1
    assert not cov.mixed()
    assert tuple(cov.coverage) == tuple([1] * 8)
    assert 1 in cov.mutations
    cov.deserialize(cov.serialize())
    assert tuple(cov.coverage) == tuple([1] * 8)
    assert 1 in cov.mutations
    cov.apply_hit(Hit(seq_nr=0, file_pos=-1, seq_pos=-2, length=8, readlength=10), 'AACCGGTT', on_plus_strand=True)
    cov.apply_hit(Hit(seq_nr=0, file_pos=-1, seq_pos=-1, length=8, readlength=10), 'ATCCGGTTA', on_plus_strand=True)
    assert 0.65 < cov.minf() < 0.69
    assert cov.mixed()
    fs = cov.fractions_at(1)
    assert list(fs.keys())[0] == 'T' and list(fs.values())[0] > 0.65
    assert list(fs.keys())[1] == 'A' and list(fs.values())[1] < 0.35


def test_spoligo_code():
    """testsuites/MTBC/spoligo.py:12-31; '400000000000001' is the answer tests/test_analyser.py:64 pins"""
    assert analyse.spoligo_code([0, 42]) == '400000000000001'
    assert analyse.spoligo_code(range(43)) == '777777777777771'
    assert analyse.spoligo_code([]) == '000000000000000'


def _plus_sequences(ts):
    out = []
    for s, k in zip(ts.plus, ts.kinds):
        m = 0 if k == 'spacer' else synth.SPACING
        out.append(Sequence(s, left=m, right=m))
    return out


def _case(tmp_path, nreads=6000):
    g = synth.genome(60000, seed=71)
    ts = synth.mtbc_templates(g, seed=72, n_phylo=10, n_res=12, regions=(3, 81), n_spacers=43)
    iso = synth.isolate(g, ts, fraction=0.5, seed=73)
    fn = str(tmp_path / 'calls.fastq')
    open(fn, 'wb').write(synth.fastq_reads(iso, nreads, readlen=150, seed=74).tobytes())
    return ts, fn


def test_fused_form_equals_apply_hit(tmp_path):
    """Coverage.from_fused over arrays derived from the hits == apply_hit over the hits (CPU only)"""
    from oracle import checker
    from util import coverage_from_hits
    if not checker.available('oracle'):
        checker.build()
    ts, fn = _case(tmp_path, 3000)
    want = checker.findseqs(fn, ts.seqs(), nthreads=1, impl='oracle', **DEFAULT)
    seqs = _plus_sequences(ts)
    a = analyse.coverages_from_hits(seqs, want['hits'], want['hitseqs'])
    cov, mut = coverage_from_hits(want['hits'], want['hitseqs'], ts.seqs(), ts.nplus)
    b = analyse.coverages_from_fused(seqs, cov, mut)
    assert [c.serialize() for c in a] == [c.serialize() for c in b]
    ca, cb = analyse.calls(ts, a), analyse.calls(ts, b)
    assert ca == cb
    assert any(ca['snps'].values()) and not all(ca['snps'].values())
    assert ca['spoligo'].startswith('7777')


@pytest.mark.gpu
def test_calls_from_gpu_equal_calls_from_reference_hits(tmp_path, checker):
    """north_star: final calls bit-exact -- the GPU's fused coverage through the restated
    consumer gives the calls the reference engine's hits give."""
    from kvarq_b200.engine import Engine
    ts, fn = _case(tmp_path)
    e = Engine(0)
    e.config(**DEFAULT)
    got = e.findseqs(fn, ts.seqs(), nplus=ts.nplus)
    impl = 'ref' if checker.available('ref') else 'oracle'
    want = checker.findseqs(fn, ts.seqs(), nthreads=1, impl=impl, **DEFAULT)
    seqs = _plus_sequences(ts)
    ref_cov = analyse.coverages_from_hits(seqs, want['hits'], want['hitseqs'])
    gpu_cov = analyse.coverages_from_fused(seqs, got.coverage, got.mutcount)
    gpu_cov_hits = analyse.coverages_from_hits(seqs, got['hits'], got['hitseqs'])
    assert [c.serialize() for c in gpu_cov] == [c.serialize() for c in ref_cov]
    assert [c.serialize() for c in gpu_cov_hits] == [c.serialize() for c in ref_cov]
    assert analyse.calls(ts, gpu_cov) == analyse.calls(ts, ref_cov)
    e.close()


# ------------------------------------------------------------------ result file
def _named(ts):
    from collections import OrderedDict
    return OrderedDict(('%s%03d' % (k, i), s) for i, (k, s) in enumerate(zip(ts.kinds, _plus_sequences(ts))))


def test_json_dump_layout():
    """kvarq/util.py:272-294: two levels indented, everything deeper on one line; still valid JSON"""
    import io
    import json
    data = {'info': {'fastq': ['a.fastq', 'b.fastq'], 'config': {'Amin': '.', 'maxerrors': 2}},
            'coverages': [['t0', '1-2-3 1[AC]'], ['t1', '0-0 ']], 'hits': [[0, 10, -3, 25, 150]]}
    out = io.StringIO()
    analyse.json_dump(data, out)
    text = out.getvalue()
    assert json.loads(text) == data
    lines = [l.rstrip(' ') for l in text.split('\n')]
    assert '    "fastq": ["a.fastq", "b.fastq"],' in lines
    assert '    ["t0", "1-2-3 1[AC]"],' in lines
    assert '    [0, 10, -3, 25, 150]' in lines
    assert max(len(l) - len(l.lstrip(' ')) for l in lines) == 4


def test_result_file_round_trip(tmp_path):
    """encode (kvarq/analyse.py:398-435) -> json_dump -> json.load -> decode gives the coverages
    back; with hits stored, update_coverages rebuilds them from the hit list (CPU only)"""
    import io
    import json
    from oracle import checker
    from kvarq_b200.fastq import Fastq
    if not checker.available('oracle'):
        checker.build()
    ts, fn = _case(tmp_path, 2000)
    want = checker.findseqs(fn, ts.seqs(), nthreads=1, impl='oracle', **DEFAULT)
    named = _named(ts)
    a = analyse.Analyser()
    fq = Fastq(fn, quiet=True)
    a.fastq_filenames, a.fastq_sizes = fq.filenames(), fq.filesizes()
    a.fastq_readlength, a.fastq_records_approx = fq.readlength, fq.records_approx
    a.config = dict(DEFAULT, nthreads=1, Azero='!')
    a.stats = {'records_parsed': 2000}
    a.hits = [analyse.Hit(*h) for h in want['hits']]
    a.hitseqs = [h.decode('latin-1') if isinstance(h, bytes) else h for h in want['hitseqs']]
    from collections import OrderedDict
    a.coverages = OrderedDict((n, Coverage(s)) for n, s in named.items())
    a.update_coverages()
    a.results = {'calls': analyse.calls(ts, list(a.coverages.values()))['spoligo']}
    out = io.StringIO()
    analyse.json_dump(a.encode(hits=True), out)
    data = json.loads(out.getvalue())
    assert data['info']['format'] == 'kvarq' and data['info']['size'] == [os.path.getsize(fn)]
    b = analyse.Analyser()
    b.decode(named, data)
    assert [c.serialize() for c in b.coverages.values()] == [c.serialize() for c in a.coverages.values()]
    assert b.hits == a.hits and b.results == a.results
    for c in b.coverages.values():                       # rebuild from the stored hits
        c.coverage = [0] * len(c.coverage)
        c.mutations = {}
    b.update_coverages()
    assert [c.serialize() for c in b.coverages.values()] == [c.serialize() for c in a.coverages.values()]


@pytest.mark.gpu
def test_scan_to_result_file(tmp_path, checker):
    """Fastq pre-flight -> engine.config -> Analyser.scan (fused arrays) -> result file: the
    coverage strings equal the ones the reference's hits give"""
    import io
    import json
    from kvarq_b200.engine import Engine
    from kvarq_b200.fastq import Fastq
    ts, fn = _case(tmp_path)
    fq = Fastq(fn, paired=True, quiet=True)
    e = Engine(0)
    e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin=fq.Q2A(13), Azero=fq.Azero)
    named = _named(ts)
    a = analyse.Analyser()
    a.scan(fq, named, engine=e)
    out = io.StringIO()
    analyse.json_dump(a.encode(), out)
    data = json.loads(out.getvalue())
    impl = 'ref' if checker.available('ref') else 'oracle'
    want = checker.findseqs(fn, ts.seqs(), nthreads=1, impl=impl, **DEFAULT)
    ref_cov = analyse.coverages_from_hits(_plus_sequences(ts), want['hits'], want['hitseqs'])
    assert [s for _, s in data['coverages']] == [c.serialize() for c in ref_cov]
    assert data['info']['config']['Amin'] == '.' and data['stats']['records_parsed'] == 6000
    hits_way = analyse.Analyser()
    hits_way.scan(fq, named, engine=e, fused=False)
    assert [c.serialize() for c in hits_way.coverages.values()] == [c.serialize() for c in ref_cov]
    e.close()
