"""Calls parity on real templates (north star "final lineage / resistance calls"):
reads are simulated from genomes with planted genotypes, scanned on the GPU, and the testsuites'
calls from the kernel's fused coverage must equal the calls from the reference engine's hits
(applied one by one like kvarq/analyse.py:57-78) -- and must name the planted genotype.

  * ebola-sierraleone14 on its bundled 19 kb genome (testsuites/ebola-sierraleone14/sierraleone14.py)
  * CanSNPer francisella / coxiella at their real positions (testsuites/CanSNPer/*_snp.txt, genomes
    shipped as +-260-base windows around the SNPs: tests/golden/testsuite_data.json)
  * MTBC phylo + resistance + spoligo at their real positions on the synthetic ancestor, through
    `python -m kvarq_b200.scan` (the reference's `kvarq scan` wiring, kvarq/cli.py:57-196)
"""
import json
import os

import numpy as np
import pytest

from kvarq_b200 import analyse, genes, synth, testsuites

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
DEFAULT = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')


@pytest.fixture(scope='module')
def data():
    return json.load(open(os.path.join(GOLDEN, 'testsuite_data.json')))


def _calls_both_ways(checker, suites, fn, spacing=25):
    """(calls from the GPU's fused coverage, calls from the CPU checker's hits), plus the analyser"""
    from kvarq_b200.engine import Engine
    from kvarq_b200.fastq import Fastq
    e = Engine(0)
    fq = Fastq(fn, quiet=True)
    e.config(nthreads=4, maxerrors=2, minoverlap=25, minreadlength=25, Amin=fq.Q2A(13), Azero=fq.Azero)
    assert e.get_config()['Amin'] == '.'
    a = analyse.Analyser(spacing=spacing)
    a.scan(fq, suites, engine=e)
    a.update_testsuites()
    e.close()
    seqs = genes.template_sequences(suites, spacing)
    plus = list(seqs.values())
    impl = 'ref' if checker.available('ref') else 'oracle'
    want = checker.findseqs(fn, [s.bases for s in plus] + [s.reverse().bases for s in plus], nthreads=1, impl=impl, **DEFAULT)
    b = analyse.Analyser(spacing=spacing)
    b.testsuites = suites
    b.coverages = analyse.OrderedDict(zip(seqs.keys(), analyse.coverages_from_hits(plus, want['hits'], want['hitseqs'])))
    b.update_testsuites()
    assert [c.serialize() for c in a.coverages.values()] == [c.serialize() for c in b.coverages.values()]
    assert not any(str(v).startswith('ERROR') for v in a.results.values()), a.results
    return a.results, b.results, a


def test_ebola_calls(checker, data, tmp_path):
    eb = data['ebola']
    g = np.frombuffer(eb['genome'].encode(), dtype=np.uint8)
    planted = [(pos, base) for pos, orig, base, geno in eb['snps'] if geno == 'SL1'] + \
              [(pos, base) for pos, orig, base, geno in eb['snps'] if geno == 'SL2'][:1]
    iso = synth.plant(g, planted)
    fn = str(tmp_path / 'ebola.fastq')
    open(fn, 'wb').write(synth.fastq_reads(iso, 6000, readlen=100, seed=5).tobytes())
    suite = testsuites.count_genotype(genes.Genome.from_bases(eb['genome'], eb['identifier']), eb['snps'])
    got, want, _ = _calls_both_ways(checker, {'sierraleone14': suite}, fn)
    assert got == want
    assert sorted(got['sierraleone14']) == ['SL1:4/4', 'SL2:1/4', 'SL3:0/1']


@pytest.mark.parametrize('organism,path', [
    ('francisella_tularensis', 'T/N.1;T.1;B.1;B.2;B.3;B.5;B.6;B.7'),
    ('coxiella_burnetii', 'Root;C.7;C.9;C.10')])
def test_cansnper_calls(checker, data, tmp_path, organism, path):
    d = data['cansnper'][organism]
    derived = set(path.split(';')[1:])
    rng = np.random.default_rng(7)
    # the isolate: derived base at the SNPs of the planted path, ancestral base at every other SNP
    allele = {}
    for name, _, _, strain, pos, new, old in d['snps']:
        if name in derived:
            allele[(strain, pos)] = new if new != '/' else [b for b in 'ACGT' if b != old][0]
        else:
            allele[(strain, pos)] = old
    chunks = []
    for strain, w in d['windows'].items():
        for start, bases in w['windows'].items():
            start = int(start)
            arr = np.frombuffer(bases.encode(), dtype=np.uint8).copy()
            for (s, pos), b in allele.items():
                if s == strain and start <= pos < start + len(arr):
                    arr[pos - start] = ord(b)
            chunks.append(synth.fastq_reads(arr, 120, readlen=120, seed=int(rng.integers(1 << 30)), first_index=1000 * len(chunks)))
    fn = str(tmp_path / (organism + '.fastq'))
    open(fn, 'wb').write(np.concatenate(chunks).tobytes())
    gs = dict((s, genes.SparseGenome(w['windows'], s, size=w['size'])) for s, w in d['windows'].items())
    suite = testsuites.snp_tree(gs, d['snps'], d['tree'])
    got, want, a = _calls_both_ways(checker, {'CanSNPer/' + organism: suite}, fn)
    assert got == want
    assert got['CanSNPer/' + organism] == [path]
    assert len(a.coverages) == len(suite.tests) and a.stats['records_parsed'] == 120 * len(chunks)


def test_mtbc_calls_through_the_cli(checker, tmp_path):
    """`python -m kvarq_b200.scan -l MTBC --genome synthetic`: lineage 2 / Beijing with rpsL K43R-position and
    embB M306-position SNPs and a substitution inside the RRDR; the result file's calls == calls from the
    reference engine's hits"""
    from kvarq_b200 import scan as cli
    from kvarq_b200 import testsuite_data as D
    g = synth.mtbc_genome()
    G = genes.Genome.from_bases(g, 'MTB ancestor (synthetic)')
    suites = testsuites.mtbc(G)
    lin2 = [(pos, base) for pos, base, var in D.PHYLO_SNPS if var in ('lineage2', 'lineage_beijing')]
    rrdr_old = G.read(761155, 1)
    planted = lin2 + [(781687, 'G'), (4247431, 'A'), (761155, 'T' if rrdr_old != 'T' else 'G')]
    iso = synth.plant(g, planted)
    sites = []
    for suite in list(suites.values())[:2]:
        for t in suite.tests:
            sites.append((t.template.start - 25, t.template.stop - t.template.start + 51))
    fq = [synth.reads_around(iso, sites, per_site=70, readlen=150, seed=11)]
    # spacers 0 and 42 only (the octal code of tests/test_analyser.py:64)
    sp = np.frombuffer((('ACGT' * 30) + D.SPACERS[0] + ('TTGCA' * 30) + D.SPACERS[42] + ('GGATC' * 30)).encode(), dtype=np.uint8)
    fq.append(synth.fastq_reads(sp, 1500, readlen=100, seed=12, first_index=500000))
    fn = str(tmp_path / 'iso.fastq')
    open(fn, 'wb').write(np.concatenate(fq).tobytes())
    out = str(tmp_path / 'iso.json')
    args = cli.parser().parse_args(['-l', 'MTBC', '--genome', 'synthetic', '-t', '4', '--poll', '0.05', '-H', fn, out])
    a = cli.scan(args, out=open(os.devnull, 'w'))
    res = json.load(open(out))
    assert res['info']['config'] == {'maxerrors': 2, 'minoverlap': 25, 'minreadlength': 25, 'nthreads': 4, 'Amin': '.', 'Azero': '!'}
    assert res['info']['testsuites'] == {'MTBC/phylo': '0.7', 'MTBC/resistance': '2.0', 'MTBC/spoligo': '0.2'}
    assert len(res['coverages']) == 132 and len(res['hits']) == len(a.hits) > 1000
    with pytest.raises(SystemExit) as ex:                # will not overwrite without -f (cli.py:90-92)
        cli.scan(args, out=open(os.devnull, 'w'))
    assert ex.value.code == cli.ERROR_FILE_EXISTS
    got, want, _ = _calls_both_ways(checker, suites, fn)
    assert res['analyses'] == json.loads(json.dumps(got)) == json.loads(json.dumps(want))
    assert got['MTBC/phylo'] == 'lineage 2/beijing sublineage'
    rs = got['MTBC/resistance']
    assert any(x.startswith('Streptomycin resistance::SNP781687AG=rpsL.') for x in rs)
    assert any(x.startswith('Ethambutol resistance::SNP4247431GA=embB.') for x in rs)
    assert not any('remark' in x for x in rs)
    assert got['MTBC/spoligo'].startswith('400000000000001 ')
