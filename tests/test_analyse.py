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
