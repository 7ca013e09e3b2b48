"""The testsuites whose `_analyse` turns coverages into the final calls, restated
in Python 3 over the tables of kvarq_b200/testsuite_data.py / caller-supplied tables:

  phylo(genome)        testsuites/MTBC/phylo.py:10-93        lineage / sublineage from 27 SNPs
  resistance(genome)   testsuites/MTBC/resistance.py:10-84   58 resistance SNPs + 4 regions
  spoligo()            testsuites/MTBC/spoligo.py:34-55      43 spacers -> octal spoligotype
  snp_tree(...)        testsuites/CanSNPer/_CanSNPer.py:44-145   hierarchical SNP typing (CanSNPer)
  count_genotype(...)  testsuites/ebola-sierraleone14/sierraleone14.py:36-58
  barcode(...)         testsuites/MTBC-SNP-barcodes/coll14.py:9-30 (generic Testsuite over a SNP table)

`genome` is a kvarq_b200.genes.Genome: the reference ships no MTB ancestor genome
(`.MISSING_LARGE_BLOBS`), so the MTBC suites take whatever genome the caller has -- the real
`.bases` file, or the synthetic stand-in of kvarq_b200.synth.mtbc_genome().

Where the original's output order follows Python-2 dict iteration (lineages joined by ' // ',
sublineages by '-', CountGenotype's list) it is made deterministic: table order.
"""
from collections import OrderedDict

from . import testsuite_data as D
from .genes import (Gene, Genotype, Reference, SNP, StaticTemplate, TemplateFromGenome, Test, Testsuite)


# ------------------------------------------------------------------------------- MTBC / phylo
class Lineage(Genotype):
    def __init__(self, name, parent=None, comutant=None):
        Genotype.__init__(self, name)
        self.name, self.parent, self.comutant, self.children = name, parent, comutant, []
        if parent is not None:
            parent.children.append(self)


class PhyloTestsuite(Testsuite):
    def __init__(self, tests, roots, version):
        Testsuite.__init__(self, tests, version)
        self.roots = roots

    def __str__(self):
        return 'TB lineage SNPs'

    def score_SNPs(self, genotypes, coverages):
        """{genotype: [SNP found?, ...]} for the tests of the given genotypes (phylo.py:15-36)"""
        out = OrderedDict()
        for test in self.tests:
            if test.genotype in genotypes:
                out.setdefault(test.genotype, []).append(bool(test.template.validate(coverages[test])))
        return out

    def _analyse(self, coverages):
        mls = []
        for ml, xs in self.score_SNPs(self.roots, coverages).items():
            if sum(xs) <= 1:                        # at least two positive SNPs
                continue
            mls.append(ml.name)
            if ml.children:
                byname, comut = OrderedDict(), {}
                for sl, xs_ in self.score_SNPs(ml.children, coverages).items():
                    byname.setdefault(sl.name, []).extend(xs_)
                    if sl.comutant:
                        comut.setdefault(sl.name, []).extend([sl.comutant] * sum(xs_))
                sls = [name + ('_' + ''.join(comut[name]) if comut.get(name) else '')
                       for name, xs_ in byname.items() if sum(xs_) > 1]
                if sls:
                    mls[-1] += '/' + '-'.join(sls)
        depths = sorted(c.mean(include_margins=False) for c in coverages.values())
        remark = ''
        if depths[len(depths) // 2] < 10:
            remark += ' -- low coverage (median below 10x)'
        if sum(c.mixed() for c in coverages.values()):
            remark += ' -- mixed coverage'
        return (' // '.join(mls) if mls else '?') + remark


def phylo(genome, force=False):
    stucki12 = Reference('Stucki et al. PLoS ONE 2012')
    lin = {}
    for var, name, parent in D.LINEAGES:
        lin[var] = Lineage(name, lin[parent] if parent else None)
    tests = [Test(SNP(genome=genome, pos=pos, base=base, force=force), lin[var], stucki12) for pos, base, var in D.PHYLO_SNPS]
    return PhyloTestsuite(tests, [l for l in lin.values() if l.parent is None], '0.7')


# -------------------------------------------------------------------------- MTBC / resistance
class DrugResistance(Genotype):
    def __init__(self, drug, gene, remarks=None):
        Genotype.__init__(self, drug + ' resistance' + (' (' + remarks + ')' if remarks else ''), gene)
        self.drug, self.remarks = drug, remarks


class ResistanceTestsuite(Testsuite):
    def __init__(self, tests, rrdr, version):
        Testsuite.__init__(self, tests, version)
        self.rrdr = rrdr

    def __str__(self):
        return 'TB resistance mutations & mutation regions'

    def _analyse(self, coverages):
        out, mixed = [], False
        for test in self.tests:
            cov, tpl = coverages[test], test.template
            if isinstance(tpl, SNP):                                    # a) SNPs (resistance.py:27-40)
                minf = cov.minf()
                mixed = mixed or minf < 0.9
                if tpl.validate(cov):
                    s = str(test) + '=' + test.genotype.gene.mut2str(tpl.start, tpl.base)
                    if minf < 0.9:
                        s += ' (%d%%)' % int(100 * minf)
                    out.append(s)
                continue
            if not mixed and cov.minf() < 0.9:                          # b) regions (:43-74)
                mixed = True
            seq = tpl.seq()
            mutations = tpl.mutations(cov)
            words = []
            for pos, newbase in mutations:
                w = '%d%s%s' % (pos + tpl.start, seq[pos], newbase)
                if test.genotype.gene:
                    w += '=' + test.genotype.gene.mut2str(pos + tpl.start, newbase)
                # (the original indexes fractions_at with the position relative to the template start,
                #  i.e. a flank base, and prints the SNP loop's stale `minf`: restated as written)
                if list(cov.fractions_at(pos).values())[0] < 0.9:
                    w += ' (%d%%)' % int(100 * self._stale_minf(coverages, test))
                words.append(w)
            if test.genotype.gene.coding and tpl.transcribe() == tpl.transcribe(mutations):
                continue                                                # no amino-acid change
            if tpl.poslist and not [m for m in mutations if m[0] + tpl.start in tpl.poslist]:
                words.append('[NONE OF MUTATIONS DOCUMENTED IN REFERENCE]')
            out.append(str(test.genotype) + ' [' + ' '.join(words) + ']')
        rrdr = [t for t in self.tests if t.genotype is self.rrdr]
        assert len(rrdr) == 1
        if coverages[rrdr[0]].mean(include_margins=False) < 10:
            out.append('remark: low coverage (RRDR below 10x)')
        if mixed:
            out.append('remark: mixed coverage')
        return out

    def _stale_minf(self, coverages, region_test):
        """`minf` as the original's region branch sees it: the value left behind by the last SNP test in
        front of this region in `self.tests` (resistance.py:28,63)."""
        last = None
        for t in self.tests:
            if t is region_test:
                break
            if isinstance(t.template, SNP):
                last = t
        return coverages[last].minf() if last is not None else 0.0


def resistance(genome, force=False):
    refs = {}
    geno = OrderedDict()
    for var, drug, gname, start, stop, promoter_end, plus, remarks in D.RESISTANCE_GENES:
        geno[var] = DrugResistance(drug, Gene(genome, gname, start, stop, promoter_end=promoter_end, plus_strand=plus), remarks)
    ref = refs.setdefault('tbdream', Reference('TBDReamDB'))
    tests = [Test(SNP(genome=genome, pos=pos, orig=orig, base=base, force=force), geno[var], ref)
             for pos, orig, base, var in D.RESISTANCE_SNPS]
    for start, stop, direction, aa_pos0, poslist, var in D.RESISTANCE_REGIONS:
        tests.append(Test(TemplateFromGenome(genome, start, stop, direction=direction, aa_pos0=aa_pos0, poslist=poslist),
                          geno[var], ref))
    return ResistanceTestsuite(tests, geno['RRDR'], '2.0')


# ----------------------------------------------------------------------------- MTBC / spoligo
class SpoligoTestsuite(Testsuite):
    def __str__(self):
        return 'TB spoligos'

    def _analyse(self, coverages):
        from .analyse import spoligo_call
        return spoligo_call([coverages[t] for t in self.tests])


def spoligo():
    ref = Reference('SpolPred.cpp by Francesc Coll')
    return SpoligoTestsuite([Test(StaticTemplate(s), Genotype('spoligo%d' % i), ref) for i, s in enumerate(D.SPACERS)], '0.2')


def mtbc(genome, force=False):
    """The three MTBC suites under the names the reference's loader gives them ('MTBC/phylo', ...)."""
    return OrderedDict([('MTBC/phylo', phylo(genome, force)), ('MTBC/resistance', resistance(genome, force)),
                        ('MTBC/spoligo', spoligo())])


# ------------------------------------------------------------------- CanSNPer: SNP-tree typing
class SNPNode(object):
    """node of the typing tree: reached when ALL its SNP tests validate (_CanSNPer.py:11-42)"""

    def __init__(self, parent, name, tests):
        self.parent, self.name, self.tests, self.children = parent, name, tests, []
        if parent is not None:
            parent.children.append(self)

    def get(self, name):
        for c in self.children:
            if c.name == name:
                return c
        return None

    def path(self):
        return self.name if self.parent is None else self.parent.path() + ';' + self.name

    def is_ancestor(self, candidate):
        """is `candidate` an ancestor of this node?"""
        node = self.parent
        while node is not None:
            if node is candidate:
                return True
            node = node.parent
        return False


class SNPTreeTestsuite(Testsuite):
    def __init__(self, tests, root, version):
        Testsuite.__init__(self, tests, version)
        self.root = root

    def walk_tree(self, found, node, coverages):
        """appends the nodes whose tests all validate, leaves first (:50-66)"""
        for child in node.children:
            n = sum(1 for t in child.tests if t.template.validate(coverages[t]))
            if n == len(child.tests):
                self.walk_tree(found, child, coverages)
                found.append(child)

    def _analyse(self, coverages):
        nodes = []
        self.walk_tree(nodes, self.root, coverages)
        i = 0
        while i < len(nodes):                       # keep the longest of every chain of ancestors (:71-81)
            j = i + 1
            while j < len(nodes):
                if nodes[i].is_ancestor(nodes[j]):
                    del nodes[j]
                else:
                    j += 1
            i += 1
        return [n.path() for n in nodes]


def snp_tree(genomes, snp_rows, tree_lines, version='1.0'):
    """make_CanSNPer_testsuite (:84-145).  genomes: {strain: Genome}; snp_rows: the columns of a CanSNPer
    `*_snp.txt` (name, organism, reference, strain, position, derived base, ancestral base; '/' = any other
    base); tree_lines: the lines of the `*_tree.txt` (root first, then ';'-joined paths)."""
    tests, refs, organism = OrderedDict(), {}, None
    for name, org, refname, strain, pos, newbase, oldbase in snp_rows:
        assert name not in tests, 'SNP "%s" defined twice' % name
        organism = organism or org
        assert organism == org, 'all SNPs must be of same organism'
        ref = refs.setdefault(refname, Reference(refname))
        newbases = sorted(set('ACGT') - set(oldbase)) if newbase == '/' else sorted(set(newbase))
        tests[name] = [Test(SNP(genome=genomes[strain], pos=int(pos), base=b, orig=oldbase, force=True), Genotype(name), ref)
                       for b in newbases]
    root = None
    for line in tree_lines:
        line = line.rstrip('\n\r')
        if root is None:
            assert ';' not in line, 'first line must be root'
            root = SNPNode(None, line, [])
            continue
        node = root
        for part in line.split(';')[1:]:
            assert part in tests, 'SNP "%s" not found' % part
            node = node.get(part) or SNPNode(node, part, tests[part])
    return SNPTreeTestsuite([t for ts in tests.values() for t in ts], root, version)


# ------------------------------------------------------------------- counting SNPs by genotype
class CountGenotype(Testsuite):
    def __str__(self):
        return 'Counting SNPs by genotype'

    def _analyse(self, coverages):
        found, total = OrderedDict(), OrderedDict()
        for test in self.tests:
            if isinstance(test.template, SNP):
                ident = test.genotype.identifier
                total[ident] = total.get(ident, 0) + 1
                found[ident] = found.get(ident, 0) + (1 if test.template.validate(coverages[test]) else 0)
        return ['%s:%d/%d' % (ident, found[ident], total[ident]) for ident in total]


def count_genotype(genome, snps, reference='', version='0.2'):
    """snps: (position, original base, mutant base, genotype name) rows (sierraleone14.py:22-32)"""
    ref = Reference(reference)
    geno = {}
    return CountGenotype([Test(SNP(genome=genome, pos=pos, orig=orig, base=base), geno.setdefault(g, Genotype(g)), ref)
                          for pos, orig, base, g in snps], version)


def barcode(genome, rows, reference='', version='0.2', force=False):
    """generic Testsuite over (name, position, original base, mutant base) rows (coll14.py:9-30)"""
    ref = Reference(reference)
    return Testsuite([Test(SNP(genome=genome, pos=pos, orig=orig, base=base, force=force), Genotype(name), ref)
                      for name, pos, orig, base in rows], version)
