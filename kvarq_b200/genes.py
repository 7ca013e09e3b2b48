"""Genetics model behind the calls: a Python-3 restatement of the parts of
kvarq/genes.py that turn coverages into genotype strings -- `Genome` (:22-108), `Gene.mut2str`
(:111-190), `Sequence` (:193-311), the templates (`StaticTemplate`, `TemplateFromGenome`, `SNP`,
:314-508) and the generic `Testsuite` (:560-632).  Host-side call logic over a few hundred
coverages; nothing here is on the GPU path.

Differences from the original, on purpose:
  * a `Genome` can be built from bases in memory (`Genome.from_bases`) or from sparse windows
    (`SparseGenome`): the test fixtures ship the neighbourhoods of the SNPs, not whole genomes;
  * Python-2 integer division is written `//` (codon numbers, `SNP.validate`);
  * where the original's answer depends on Python-2 dict / set iteration order (ties between equally
    frequent alternate bases, the order of `'/'`-expanded CanSNPer bases) the order is made
    deterministic (by base);
  * testsuite files are not `execfile`d: kvarq_b200/testsuites.py builds the suites from tables.
"""
import re

from .analyse import PAIRS, Sequence as _BaseSequence

COMPATIBILITY = '0.2'                   # kvarq/genes.py:19

_AA = 'FFLLSSSSYY$$CC$WLLLLPPPPHHQQRRRRIIIMTTTTNNKKSSRRVVVVAAAADDEEGGGG'     # TCAG x TCAG x TCAG
CODE = dict((a + b + c, _AA[16 * i + 4 * j + k]) for i, a in enumerate('TCAG') for j, b in enumerate('TCAG')
            for k, c in enumerate('TCAG'))                                      # DNA codon table (genes.py:208-218)


class Sequence(_BaseSequence):
    """analyse.Sequence plus the codon machinery (genes.py:248-311)"""
    pairs = PAIRS
    code = CODE

    def __setitem__(self, idx, value):
        if idx < 0 or idx > len(self.bases) - 1:
            raise IndexError
        self.bases = self.bases[:idx] + value + self.bases[idx + 1:]

    def reverse(self):
        r = _BaseSequence.reverse(self)
        return Sequence(r.bases, left=r.left, right=r.right, pos=r.pos, plus_strand=r.plus_strand)

    def get_aa(self, pos, mutations=()):
        first = pos - pos % 3
        codon = list(self.bases[first:first + 3])
        for mpos, base in mutations:
            if 0 <= mpos - first < 3:
                codon[mpos - first] = base
        return self.code[''.join(codon)]

    def transcribe(self, mutations=()):
        return ''.join(self.get_aa(3 * i, mutations) for i in range(len(self) // 3))

    def apply_mutations(self, mutations):
        bases = list(self.bases)
        for pos, base in mutations:
            bases[pos] = base
        self.bases = ''.join(bases)


class Genome(object):
    """A reference genome addressed with 1-based positions (genes.py:22-108): a `.bases` file
    (raw letters) or a FASTA file (first record)."""

    def __init__(self, path, identifier=None, description=None):
        self.path = path
        with open(path, 'rb') as f:
            raw = f.read()
        if raw[:1] == b'>':
            head, _, rest = raw.partition(b'\n')
            head = head.decode('latin-1').rstrip('\r')
            name, _, descr = head[1:].partition(' ')
            identifier = identifier if identifier is not None else name
            description = description if description is not None else (descr + '\n' if descr else None)
            if b'>' in rest:
                rest = rest[:rest.index(b'>')]
            self.bases = rest.replace(b'\n', b'').replace(b'\r', b'').decode('latin-1')
            self.fasta = True
        else:
            self.bases = raw.decode('latin-1')      # (a trailing newline counts, as in the original's size)
            self.fasta = False
        self.size = len(self.bases)
        self.identifier = identifier
        self.description = description

    @classmethod
    def from_bases(cls, bases, identifier):
        g = cls.__new__(cls)
        if isinstance(bases, (bytes, bytearray)):
            bases = bytes(bases).decode('latin-1')
        elif not isinstance(bases, str):
            bases = bytes(bases).decode('latin-1')  # numpy uint8
        g.path, g.bases, g.size, g.fasta = None, bases, len(bases), True
        g.identifier, g.description = identifier, None
        return g

    def read(self, pos, length):
        return self.bases[pos - 1:pos - 1 + length]

    def seq(self, start, stop, left=0, right=0, **kw):
        return Sequence(self.read(start - left, stop - start + 1 + left + right), left, right, pos=start - left, **kw)

    def __str__(self):
        return self.identifier


class SparseGenome(Genome):
    """Only some windows of a genome are known: {first position (1-based): bases}.  Reading outside
    them is an error -- templates of a testsuite must lie inside the shipped neighbourhoods."""

    def __init__(self, windows, identifier, size=None):
        self.windows = sorted((int(p), b) for p, b in windows.items())
        self.identifier, self.description, self.path, self.fasta = identifier, None, None, True
        self.size = size if size is not None else max(p + len(b) - 1 for p, b in self.windows)

    def read(self, pos, length):
        for p, b in self.windows:
            if p <= pos and pos + length <= p + len(b):
                return b[pos - p:pos - p + length]
        raise KeyError('%s: bases %d..%d are not in the shipped windows' % (self.identifier, pos, pos + length - 1))


class Gene(object):
    """genes.py:111-190"""

    def __init__(self, genome, identifier, start, stop, promoter_end=None, plus_strand=True, coding=True):
        assert start <= stop, 'start position must be smaller than stop position'
        self.genome, self.identifier = genome, identifier
        self.start, self.stop = start, stop
        self.promoter_end = start if promoter_end is None else promoter_end
        self.plus_strand, self.coding = plus_strand, coding

    def mut2str(self, pos, newbase):
        """'gene.XnY' (amino acids; n = codon number), 'gene.nXY' for non-coding genes (bases),
        'gene promoter mutation -n' in front of the gene, '?' elsewhere.  `newbase` is on the plus strand."""
        if pos < self.promoter_end:
            return '%s promoter mutation %d' % (self.identifier, pos - self.promoter_end)
        if pos < self.start or pos > self.stop:
            return '?'
        pos1 = pos - self.start + 1
        codon_nr = (pos - self.start) // 3 + 1
        codon_start = self.start + (codon_nr - 1) * 3
        codon_mut = pos - codon_start
        codon = self.genome.seq(codon_start, codon_start + 2)
        oldbase = self.genome.read(pos, 1)
        if not self.plus_strand:
            pos1 = self.stop - pos + 1
            codon_nr = (self.stop - pos) // 3 + 1
            codon_mut = 2 - codon_mut
            codon = codon.reverse()
            newbase, oldbase = PAIRS[newbase], PAIRS[oldbase]
        if self.coding:
            return '%s.%s%d%s' % (self.identifier, codon.transcribe(), codon_nr,
                                  codon.transcribe(mutations=((codon_mut, newbase),)))
        return '%s.%d%s%s' % (self.identifier, pos1, oldbase, newbase)

    def __str__(self):
        return ('gene %s %d..%d' if self.plus_strand else 'gene %s complement(%d..%d)') % (self.identifier, self.start, self.stop)


# ------------------------------------------------------------------------------------- templates
class Template(object):
    def __init__(self, identifier):
        self.identifier = identifier

    def validate(self, coverage):                   # genes.py:326-331: "found" = mean depth without margins >= 2
        return coverage.mean(include_margins=False) >= 2

    def __str__(self):
        return self.identifier


class StaticTemplate(Template):
    def __init__(self, bases, identifier=None):
        Template.__init__(self, identifier or bases)
        self.bases = bases

    def seq(self):
        return Sequence(self.bases)


class DynamicTemplate(Template):
    """templates that can grow flanks: seq(spacing)"""


class TemplateFromGenome(DynamicTemplate):
    """genes.py:371-466"""

    def __init__(self, genome, start, stop, direction='+', aa_pos0=1, poslist=None):
        assert start <= stop and direction in '+-'
        DynamicTemplate.__init__(self, '%s[%d:%d](%s)' % (genome, start, stop, direction))
        self.genome, self.start, self.stop = genome, start, stop
        self.direction, self.aa_pos0, self.poslist = direction, aa_pos0, poslist

    def seq(self, spacing=0):
        return self.genome.seq(self.start, self.stop, spacing, spacing)

    def transcribe(self, mutations=None):
        seq = self.seq()
        if mutations:
            seq.apply_mutations(mutations)
        if self.direction == '-':
            seq = seq.reverse()
        return seq.transcribe()

    def mutations(self, coverage):
        """[(pos relative to .start, base)] of the most prevalent alternate base where it is frequent
        enough (:424-450).  Iterates positions in ascending order, ties by base."""
        out = []
        mean, std = coverage.mean(), coverage.std()
        n_seq = len(self.seq())
        for cpos in sorted(coverage.mutations):
            bases = coverage.mutations[cpos]
            if cpos < coverage.start or cpos - coverage.start >= n_seq:
                continue
            base, n = sorted(((b, bases.count(b)) for b in set(bases)), key=lambda x: (-x[1], x[0]))[0]
            if n > 1 and n > mean - 1.5 * std:
                out.append((cpos - coverage.start, base))
        return out

    def aa_mutations(self, mutations):
        before, after = self.transcribe(), self.transcribe(mutations)
        return [(i + self.aa_pos0, a, after[i]) for i, a in enumerate(before) if after[i] != a]


class SNP(TemplateFromGenome):
    """The template of a SNP is the MUTANT sequence (:469-508)."""

    def __init__(self, genome, pos, base, orig=None, force=False):
        TemplateFromGenome.__init__(self, genome, pos, pos)
        self.base, self.orig = base, orig
        oldbase = genome.read(pos, 1)
        if not force:
            if orig:
                assert oldbase == orig, 'expected orig %s found %s' % (orig, oldbase)
            assert base != oldbase
        self.identifier = 'SNP%d%s%s' % (pos, oldbase, base)

    def seq(self, spacing=0):
        seq = TemplateFromGenome.seq(self, spacing=spacing)
        seq[spacing] = self.base
        return seq

    def validate(self, coverage):
        c = coverage.coverage[coverage.start]
        m = len(coverage.mutations.get(coverage.start, ''))
        return c >= 2 and m < c // 2


# ------------------------------------------------------------------------ tests and testsuites
class Reference(object):
    def __init__(self, descr):
        self.descr = descr


class Genotype(object):
    def __init__(self, identifier, gene=None):
        self.identifier, self.gene = identifier, gene

    def __str__(self):
        return str(self.identifier)

    def __repr__(self):
        return '<%s : "%s">' % (self.__class__.__name__, self.identifier)


class Test(object):
    __test__ = False                                # (not a pytest class)

    def __init__(self, template, genotype, reference):
        assert '::' not in str(template)
        self.template, self.genotype, self.reference = template, genotype, reference

    def __str__(self):
        return '%s::%s' % (self.genotype, self.template)


class AnalysisException(RuntimeError):
    pass


class Testsuite(object):
    """genes.py:560-632: the generic suite lists validated SNPs and region mutations."""
    __test__ = False

    def __init__(self, tests, version):
        self.tests, self.version = tests, version

    def _analyse(self, coverages):
        out = []
        for test in self.tests:
            cov, tpl = coverages[test], test.template
            if isinstance(tpl, SNP):
                if tpl.validate(cov):
                    out.append(str(test))
            elif isinstance(tpl, TemplateFromGenome):
                seq = tpl.seq()
                for pos, newbase in tpl.mutations(cov):
                    s = '%d%s%s' % (pos + tpl.start, seq[pos], newbase)
                    if test.genotype.gene:
                        s += '=' + test.genotype.gene.mut2str(pos + tpl.start, newbase)
                    out.append(s)
        return out

    def analyse(self, analyser):
        try:
            coverages = dict((test, analyser[test]) for test in self.tests)
        except KeyError as e:
            raise AnalysisException('template "%s" not found' % e.args[0])
        return self._analyse(coverages)

    def __str__(self):
        return 'generic Testsuite with %d tests' % len(self.tests)


def template_sequences(testsuites, spacing):
    """Analyser.load_coverages (kvarq/analyse.py:260-282): OrderedDict str(template) -> Sequence for every
    test of every suite, flanks of `spacing` bases on templates that can grow them."""
    from collections import OrderedDict
    out = OrderedDict()
    for suite in testsuites.values():
        for test in suite.tests:
            tpl = test.template
            out[str(tpl)] = tpl.seq(spacing=spacing) if isinstance(tpl, DynamicTemplate) else tpl.seq()
    return out
