"""The genetics model and the testsuites' call logic (kvarq_b200/genes.py, testsuites.py): the
reachable cases of the reference's tests/test_genes.py (the others need the MTB ancestor genome the
reference does not ship: the same checks are made on genomes built here, with answers worked out
by hand), the CanSNPer tree-walk cases of testsuites/CanSNPer/_test.py, and the MTBC call trees on
hand-made coverages."""
import json
import os

import pytest

from kvarq_b200 import analyse, genes, synth, testsuites

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


@pytest.fixture(scope='module')
def data():
    return json.load(open(os.path.join(GOLDEN, 'testsuite_data.json')))


# ---------------------------------------------------------------- tests/test_genes.py, reachable
def test_reverse():
    assert genes.Sequence('AAACGT').reverse().bases == 'ACGTTT'                 # test_genes.py:46-47


def test_code():
    seq = genes.Sequence('GCTTGTGATTGC')                                         # test_genes.py:50-56
    for i in range(4):
        for j in range(3):
            assert seq.get_aa(i * 3 + j) == 'ACDC'[i]
    assert seq.get_aa(1, [(1, 'T')]) == 'V'
    assert seq.transcribe() == 'ACDC'
    assert len(genes.CODE) == 64 and sorted(set(genes.CODE.values())) == sorted('ACDEFGHIKLMNPQRSTVWY$')
    assert genes.CODE['ATG'] == 'M' and genes.CODE['TGA'] == '$' and genes.CODE['TGG'] == 'W' and genes.CODE['AGA'] == 'R'


def test_genome_bases_and_fasta(tmp_path):
    """test_genome (:116-127): a .bases file and a FASTA file of the same sequence read the same"""
    g = synth.genome(1000, seed=4).tobytes().decode()
    open(str(tmp_path / 'g.bases'), 'w').write(g)
    with open(str(tmp_path / 'g.fa'), 'w') as f:
        f.write('>TEST some description\n')
        for i in range(0, 1000, 70):
            f.write(g[i:i + 70] + '\n')
        f.write('>SECOND\nACGT\n')
    g1, g2 = genes.Genome(str(tmp_path / 'g.bases')), genes.Genome(str(tmp_path / 'g.fa'))
    assert g2.identifier == 'TEST' and g2.size == 1000 and g1.size == 1000
    for pos, length in ((1, 50), (17, 100), (900, 100), (951, 50)):
        assert g1.read(pos, length) == g2.read(pos, length) == g[pos - 1:pos - 1 + length]
    s = g2.seq(101, 110, 5, 7)
    assert s.bases == g[95:117] and s.left == 5 and s.right == 7 and s.pos == 96


def _toy_genome():
    #        1         11        21        31
    bases = 'CCATGACAGGTTTGCATAACCGGTTAACC' + 'GATTACA' * 10
    return genes.Genome.from_bases(bases, 'toy')


def test_gene_mut2str():
    """Gene.mut2str (test_gene, :15-33, on a toy genome): plus / minus strand, promoter, non-coding"""
    g = _toy_genome()
    fwd = genes.Gene(g, 'fwd', 3, 20)                       # ATG ACA GGT TTG CAT AAC
    assert fwd.mut2str(7, 'G') == 'fwd.T2R'                 # ACA -> AGA
    assert fwd.mut2str(3, 'G') == 'fwd.M1V'                 # ATG -> GTG
    assert fwd.mut2str(2, 'A') == 'fwd promoter mutation -1'
    assert fwd.mut2str(21, 'A') == '?'
    rev = genes.Gene(g, 'rev', 3, 20, plus_strand=False)    # read from the minus strand: GTT ATG CAA ACC TGT CAT
    assert rev.mut2str(20, 'A') == 'rev.V1F'                # plus C -> A at the last base = G -> T at the first: GTT -> TTT
    assert rev.mut2str(16, 'C') == 'rev.M2R'                # plus A -> C = T -> G in the middle of ATG: AGG
    nc = genes.Gene(g, 'rna', 3, 20, coding=False)
    assert nc.mut2str(7, 'G') == 'rna.5CG'
    ncr = genes.Gene(g, 'rna_rev', 3, 20, coding=False, plus_strand=False)
    assert ncr.mut2str(19, 'T') == 'rna_rev.2TA'            # plus A -> T reads T -> A on the minus strand


def test_snp_template_and_validate():
    """SNP.seq plants the mutant base behind the left flank (test_SNP, :35-43); SNP.validate and
    TemplateFromGenome.mutations thresholds (test_mutations, :75-113)"""
    g = _toy_genome()
    snp = genes.SNP(g, 7, base='G', orig='C')
    assert str(snp) == 'SNP7CG' and snp.seq(spacing=3).bases == 'TGAGAGG' and snp.seq(spacing=3).left == 3
    with pytest.raises(AssertionError):
        genes.SNP(g, 7, base='C')
    with pytest.raises(AssertionError):
        genes.SNP(g, 7, base='G', orig='T')
    cov = analyse.Coverage(snp.seq(spacing=3))
    assert not snp.validate(cov)                            # no coverage
    cov.mutations = {3: 'C'}
    assert not snp.validate(cov)
    cov.coverage = [20] * len(cov.coverage)
    cov.mutations = {3: 'C' * 10}
    assert not snp.validate(cov)                            # half of the reads disagree
    cov.mutations = {3: 'C'}
    assert snp.validate(cov)
    region = genes.TemplateFromGenome(g, 3, 20)
    rc = analyse.Coverage(region.seq(spacing=2))
    rc.mutations = {2 + 0: 'TGC', 2 + 4: 'GGGGGGGGGGGAT'}
    rc.coverage = [1000] * len(rc.coverage)
    assert region.mutations(rc) == []                       # not validated at very high coverage
    rc.coverage = [10] * len(rc.coverage)
    assert region.mutations(rc) == [(4, 'G')]
    assert region.aa_mutations([(4, 'G')]) == [(2, 'T', 'R')]
    assert region.transcribe() == 'MTGLHN' and genes.TemplateFromGenome(g, 3, 20, direction='-').transcribe() == 'VMQTCH'


def test_sparse_genome(data):
    w = data['cansnper']['coxiella_burnetii']['windows']['AE016828.2']
    g = genes.SparseGenome(w['windows'], 'AE016828.2', size=w['size'])
    name, _, _, strain, pos, new, old = data['cansnper']['coxiella_burnetii']['snps'][0]
    assert len(g.read(pos - 25, 51)) == 51 and g.seq(pos, pos, 25, 25).pos == pos - 25
    with pytest.raises(KeyError):
        g.read(1, 10 ** 6)


# ------------------------------------------------------------ testsuites/CanSNPer/_test.py
class _Dummy(genes.Genome):
    def __init__(self):
        self.identifier, self.size = 'dummy', 0

    def read(self, pos, length):
        return 'N' * length


def _cansnper(data, org):
    d = data['cansnper'][org]
    return testsuites.snp_tree(dict((s, _Dummy()) for s in set(r[3] for r in d['snps'])), d['snps'], d['tree'])


def _fake_coverages(suite, names):
    for t in suite.tests:
        t.template.validate = lambda x: x
    return dict((t, t.genotype.identifier in names) for t in suite.tests)


def test_cansnper_tree_walk(data):
    ba = _cansnper(data, 'bacillus_anthracis')
    ret = ba._analyse(_fake_coverages(ba, 'A/B.Br.001;B.Br.003'.split(';')))
    assert 'A/B.Br.001;B.Br.003' in ret and 'A/B.Br.001;B.Br.003;B.Br.004' not in ret and 'A/B.Br.001' not in ret
    ret = ba._analyse(_fake_coverages(ba, 'A/B.Br.001;A.Br.006;A.Br.004;A.Br.003;A_Br_002;A.Br.001'.split(';')))
    assert ret == ['A/B.Br.001;A.Br.006;A.Br.004;A.Br.003']                    # a node is missing: the walk stops there
    yp = _cansnper(data, 'yersinia_pestis')
    path = 'Root;0.PE3.a;III;VI;VII;0.ANT3.a;3.ANT.a;VIII;2.MED3.a;IX;X;2.MED1.a;2.MED1.b;2.MED1.d'
    assert yp._analyse(_fake_coverages(yp, path.split(';'))) == [path]
    cb = _cansnper(data, 'coxiella_burnetii')
    assert cb._analyse(_fake_coverages(cb, 'Root;C.7;C.9;C.10'.split(';'))) == ['Root;C.7;C.9;C.10']
    assert sum(len(data['cansnper'][o]['snps']) for o in data['cansnper']) == 188


# ----------------------------------------------------------------------- MTBC call trees
def _coverages(suite, present, depth=30, spacing=25):
    """hand-made coverages: every template covered `depth` times; SNP templates not in `present` carry the
    original base at the SNP position in every read (= a mutation against the mutant template)"""
    covs = {}
    for t in suite.tests:
        tpl = t.template
        seq = tpl.seq(spacing=spacing) if isinstance(tpl, genes.DynamicTemplate) else tpl.seq()
        c = analyse.Coverage(seq)
        c.coverage = [depth] * len(seq)
        if isinstance(tpl, genes.SNP) and str(tpl) not in present:
            c.mutations = {c.start: tpl.genome.read(tpl.start, 1) * depth}
        covs[t] = c
    return covs


def test_mtbc_call_trees():
    G = genes.Genome.from_bases(synth.mtbc_genome(), 'MTB ancestor (synthetic)')
    suites = testsuites.mtbc(G)
    assert len(genes.template_sequences(suites, 25)) == 132 and sum(len(s) for s in genes.template_sequences(suites, 25).values()) == 6318
    ph = suites['MTBC/phylo']
    l2 = [str(t.template) for t in ph.tests if t.genotype.name in ('lineage 2', 'beijing sublineage')]
    assert ph._analyse(_coverages(ph, l2)) == 'lineage 2/beijing sublineage'     # the string of tests/test_analyser.py:65-66
    l3 = [str(t.template) for t in ph.tests if t.genotype.name == 'lineage 3']
    assert ph._analyse(_coverages(ph, l3, depth=6)) == 'lineage 3 -- low coverage (median below 10x)'
    assert ph._analyse(_coverages(ph, l3[:1])) == '?'
    both = ph._analyse(_coverages(ph, l2[:3] + l3))
    assert both == 'lineage 2 // lineage 3'
    rs = suites['MTBC/resistance']
    assert rs._analyse(_coverages(rs, [])) == []
    snp = [t for t in rs.tests if str(t.template).startswith('SNP781687')][0]
    out = rs._analyse(_coverages(rs, [str(snp.template)]))
    assert len(out) == 1 and out[0].startswith('Streptomycin resistance::SNP781687AG=rpsL.') and out[0].endswith('43' + out[0][-1])
    low = rs._analyse(_coverages(rs, [], depth=5))
    assert low == ['remark: low coverage (RRDR below 10x)']
    # a region mutation: every read carries another base at RRDR position 761155
    covs = _coverages(rs, [])
    rrdr = [t for t in rs.tests if t.genotype is rs.rrdr][0]
    c = covs[rrdr]
    old = G.read(761155, 1)
    new = 'A' if old != 'A' else 'C'
    c.mutations = {c.start + (761155 - 761082): new * 30}
    out = rs._analyse(covs)
    codon = (761155 - 759807) // 3 + 1
    if out:                                                # (a synonymous change prints nothing)
        assert out[0].startswith('Rifampicin resistance (RRDR) [761155%s%s=rpoB.' % (old, new)) and str(codon) in out[0]
    sp = suites['MTBC/spoligo']
    covs = _coverages(sp, [])
    for i, t in enumerate(sp.tests):
        if i not in (0, 42):
            covs[t].coverage = [0] * 25
    assert sp._analyse(covs) == '400000000000001 1000000000000000000000000000000000000000001'
