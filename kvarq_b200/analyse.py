"""Python-3 restatement of the consumer of the hit list: `Coverage`
(kvarq/analyse.py:25-185), the parts of `Sequence` it needs (kvarq/genes.py:193-276), the call
primitives the MTBC testsuites are made of -- `Template.validate` (genes.py:326-331),
`SNP.validate` (:503-508), `TemplateFromGenome.mutations` (:424-450), the spoligotype code
(testsuites/MTBC/spoligo.py:12-55) -- and the hit -> coverage mapping of `Analyser`
(analyse.py:284-294,352-395).

Two ways to fill a Coverage give identical objects: `apply_hit` over the hits (the reference's
way) and `Coverage.from_fused` over the arrays the scan kernel accumulates (`coverage[]`,
`mutcount[][6]`).  The reference appends alternate bases in hit order; every consumer only counts
or sorts them, so the fused form keeps them sorted.

Python-2 traps of the original that matter here: integer `/` in `validate`
(`m < c/2`), `dict.values()[0]` in `minf`, `oct()` formatting in the spoligo code.
"""
from collections import OrderedDict, namedtuple

Hit = namedtuple('Hit', 'seq_nr file_pos seq_pos length readlength')

PAIRS = {'A': 'T', 'T': 'A', 'G': 'C', 'C': 'G', 'N': 'N'}       # Sequence.pairs, genes.py:204
COLUMNS = 'ACGTN'                                                # mutcount columns (5 = other)


class Sequence(object):
    """bases with margins `left` / `right` (genes.py:193-276, the parts Coverage touches)"""

    def __init__(self, bases, left=0, right=0, pos=None, plus_strand=True):
        if isinstance(bases, (bytes, bytearray)):
            bases = bytes(bases).decode('latin-1')
        self.bases = bases
        self.left = left
        self.right = right
        self.pos = pos
        self.plus_strand = plus_strand

    def __len__(self):
        return len(self.bases)

    def __getitem__(self, idx):
        return self.bases[idx]

    def reverse(self):
        return Sequence(''.join([PAIRS[b] for b in self.bases])[::-1], pos=self.pos,
                        plus_strand=not self.plus_strand, left=self.left, right=self.right)

    def plus_idx(self, idx):
        return idx if self.plus_strand else len(self.bases) - idx - 1

    def plus_base(self, base):
        return base if self.plus_strand else PAIRS[base]


class Coverage(object):
    """kvarq/analyse.py:25-185"""

    def __init__(self, plus_seq):
        self.plus_seq = plus_seq
        self.minus_seq = plus_seq.reverse()
        self.coverage = [0] * len(plus_seq)
        self.mutations = {}
        self.start = plus_seq.left
        self.stop = len(plus_seq) - plus_seq.right

    @classmethod
    def from_fused(cls, plus_seq, coverage, mutcount):
        """coverage[len], mutcount[len][6] (A,C,G,T,N,other on the plus strand) of one template,
        as accumulated by the scan kernel (kvq_result_coverage / kvq_result_mutcount)."""
        c = cls(plus_seq)
        c.coverage = [int(x) for x in coverage]
        for idx, row in enumerate(mutcount):
            if int(row[5]):
                raise ValueError('alternate bases outside ACGTN at %d: the reference would raise KeyError '
                                 'on the minus strand' % idx)
            s = ''.join(COLUMNS[k] * int(row[k]) for k in range(5))
            if s:
                c.mutations[idx] = s
        return c

    def apply_hit(self, hit, hitseq, on_plus_strand):                  # :57-78
        seq = self.plus_seq if on_plus_strand else self.minus_seq
        start = max(0, hit.seq_pos)
        stop = start + hit.length
        for i, j in enumerate(range(start, stop)):
            c_j = seq.plus_idx(j)
            c_b = seq.plus_base(hitseq[i])
            self.coverage[c_j] += 1
            if hitseq[i] != seq[j]:
                self.mutations[c_j] = self.mutations.get(c_j, '') + c_b

    def bases_at(self, idx):                                            # :80-87
        m = self.mutations.get(idx, '')
        ret = {self.plus_seq[idx]: self.coverage[idx] - len(m)}
        for b in set(m):
            ret[b] = m.count(b)
        return ret

    def fractions_at(self, idx):                                        # :89-96
        bases = self.bases_at(idx)
        total = sum(bases.values())
        # ties: Python 2 sorted() is stable over dict order (arbitrary); break them by base
        return OrderedDict(sorted([(b, n / float(max(1, total))) for b, n in bases.items()],
                                  key=lambda x: (-x[1], x[0])))

    def minf(self, include_margins=False):                              # :98-110
        start, stop = (0, len(self)) if include_margins else (self.start, self.stop)
        return min([list(self.fractions_at(pos).values())[0] for pos in range(start, stop)])

    def mixed(self, fmin=0.9, include_margins=False):                   # :112-123
        cminf = self.minf(include_margins=include_margins)
        return cminf > 0 and cminf < fmin

    def mean(self, include_margins=True):                               # :125-132
        if include_margins:
            return sum(self.coverage) / float(len(self.coverage))
        return sum(self.coverage[self.start:self.stop]) / float(self.stop - self.start)

    def std(self, include_margins=True):                                # :134-143 (margins ignored there too)
        m = self.mean()
        return (sum([(x - m) ** 2 for x in self.coverage]) / len(self.coverage)) ** .5

    def seqmean(self):                                                  # :145-149
        seq = [self.coverage[i] - len(self.mutations.get(i, '')) for i in range(self.start, self.stop)]
        return sum(seq) / float(len(seq))

    def serialize(self):                                                # :156-164
        cov = '-'.join([str(c) for c in self.coverage])
        mut = '-'.join(['%d[%s]' % (idx, ''.join(sorted(self.mutations[idx])))
                        for idx in sorted(self.mutations.keys())])
        return cov + ' ' + mut

    def deserialize(self, serialized_coverage):                         # :166-178
        c_s, _, m_s = serialized_coverage.partition(' ')
        self.coverage = [int(x) for x in c_s.split('-')]
        if m_s:
            self.mutations = dict([(int(x[:x.index('[')]), x[x.index('[') + 1:x.index(']')])
                                   for x in m_s.split('-')])
        else:
            self.mutations = {}

    def __str__(self):
        return '(mean %.2f std %.2f) ' % (self.mean(), self.std()) + ' ' + self.serialize()

    def __len__(self):
        return len(self.coverage)

    def __getitem__(self, idx):
        return self.coverage[idx]


# ---------------------------------------------------------------- hits / fused arrays -> coverages
def coverages_from_hits(plus_seqs, hits, hitseqs):
    """Analyser.update_coverages (analyse.py:352-395): the scan ran over `plus + reversed`, so
    seq_nr >= N means the minus strand of template seq_nr - N (coverage_at, :284-294)."""
    covs = [Coverage(s) for s in plus_seqs]
    n = len(covs)
    for hit, hs in zip(hits, hitseqs):
        hit = Hit(*hit)
        if isinstance(hs, (bytes, bytearray)):
            hs = bytes(hs).decode('latin-1')
        covs[hit.seq_nr - n if hit.seq_nr >= n else hit.seq_nr].apply_hit(hit, hs, hit.seq_nr < n)
    return covs


def coverages_from_fused(plus_seqs, coverage, mutcount):
    """The same coverages from the arrays the kernel accumulated (concatenated plus templates)."""
    covs, at = [], 0
    for s in plus_seqs:
        covs.append(Coverage.from_fused(s, coverage[at:at + len(s)], mutcount[at:at + len(s)]))
        at += len(s)
    return covs


# ------------------------------------------------------------------------------ call primitives
def template_validate(coverage):
    """Template.validate (genes.py:326-331): found iff mean depth without margins >= 2"""
    return coverage.mean(include_margins=False) >= 2


def snp_validate(coverage):
    """SNP.validate (genes.py:503-508); `c/2` is Python-2 integer division"""
    c = coverage.coverage[coverage.start]
    m = len(coverage.mutations.get(coverage.start, []))
    return c >= 2 and m < c // 2


def region_mutations(coverage):
    """TemplateFromGenome.mutations (genes.py:424-450): [(pos relative to .start, base), ...] of
    the most prevalent alternate base where it is frequent enough."""
    ret = []
    mean = coverage.mean()
    std = coverage.std()
    seqlen = coverage.stop - coverage.start
    for cpos in sorted(coverage.mutations):
        bases = coverage.mutations[cpos]
        if cpos < coverage.start or cpos - coverage.start >= seqlen:
            continue
        # most prevalent alternate base; ties (dict order in Python 2) broken by base
        base, n = sorted([(b, bases.count(b)) for b in set(bases)], key=lambda x: (-x[1], x[0]))[0]
        if n > 1 and n > mean - 1.5 * std:
            ret.append((cpos - coverage.start, base))
    return ret


def spoligo_code(spnrs):
    """testsuites/MTBC/spoligo.py:12-31: 14 octal digits for spacers 0..41 + one binary digit"""
    value = 0
    spoligo42 = '0'
    for spnr in spnrs:
        if spnr == 42:
            spoligo42 = '1'
        else:
            value += 2 ** (41 - spnr)
    octal = '%o' % value
    return '0' * (14 - len(octal)) + octal + spoligo42


def spoligo_call(coverages):
    """SpoligoTestsuite._analyse (spoligo.py:39-55) over the 43 spacer coverages.  (The original
    averages `coverages[self.tests[spnr]]` with a stale loop variable; restated as written.)"""
    spnrs = [i for i, c in enumerate(coverages) if template_validate(c)]
    # `spnr` in the original's sum is the comprehension variable leaked by Python 2: the last index
    # enumerated, i.e. always the last spacer
    last = len(coverages) - 1
    if not spnrs or sum([coverages[last].mean() for _ in spnrs]) / len(spnrs) < 10:
        remark = ' -- low coverage (mean below 10x)'
    else:
        remark = ''
    spoct = spoligo_code(spnrs)
    spbin = ''.join([bin(int(x))[2:].rjust(3, '0') for x in spoct[:14]] + [bin(int(x))[2:] for x in spoct[14:]])
    return ' '.join([spoct, spbin]) + remark


def depth_remarks(coverages):
    """the remarks PhyloTestsuite._analyse appends (testsuites/MTBC/phylo.py:80-90)"""
    depths = sorted([c.mean(include_margins=False) for c in coverages])
    remark = ''
    if depths[len(depths) // 2] < 10:
        remark += ' -- low coverage (median below 10x)'
    if sum([c.mixed() for c in coverages]):
        remark += ' -- mixed coverage'
    return remark


def calls(template_set, coverages):
    """All calls for a synth.TemplateSet-shaped template list (kinds 'phylo' / 'res' SNPs,
    'region', 'spacer'): which SNPs are present, region mutations, spoligotype, remarks."""
    out = {'snps': {}, 'regions': {}, 'spoligo': None, 'remarks': ''}
    spacers, snpcov = [], []
    for i, (kind, cov) in enumerate(zip(template_set.kinds, coverages)):
        if kind in ('phylo', 'res', 'panel'):
            out['snps'][i] = snp_validate(cov)
            snpcov.append(cov)
        elif kind == 'region':
            out['regions'][i] = region_mutations(cov)
        elif kind == 'spacer':
            spacers.append(cov)
    if spacers:
        out['spoligo'] = spoligo_call(spacers)
    if snpcov:
        out['remarks'] = depth_remarks(snpcov)
    return out


# ------------------------------------------------------------------ result file
VERSION = '0.12.2-b200'             # info.version: the reference writes kvarq.VERSION here


def json_dump(data, fd, indent=2, max_indent_level=2):
    """kvarq/util.py:272-294: indent=2 JSON whose levels deeper than `max_indent_level` are folded
    onto one line (a coverage string, a hit, the readlengths tuple stay on their line).  Python 2's
    encoder separates items by ', ' also when indenting; the separators are set to match."""
    import json
    import re
    deep = indent * max_indent_level
    fold = re.compile('^([\\[{,]? ?)\n {%d,}' % (deep + 1), re.MULTILINE)      # break into a deeper level
    closing = re.compile('^\n {%d}$' % deep, re.MULTILINE)                      # break back to the last kept level
    held = ''
    for chunk in json.JSONEncoder(indent=2, separators=(', ', ': ')).iterencode(data):
        if closing.match(chunk):
            held = chunk
            continue
        if held and chunk not in ('}', ']', ','):
            fd.write(held)
        held = ''
        fd.write(fold.sub('\\1', chunk))


class Analyser(object):
    """The scan -> coverages -> result-file part of kvarq/analyse.py:208-435 over named templates
    (`OrderedDict name -> Sequence`; the reference derives them from testsuites,
    load_coverages :261-282).  `scan` asks the engine for the fused arrays, so no Python loop
    over hits is needed; `update_coverages` is the reference's loop, for results read back
    from a file with hits."""

    def __init__(self, spacing=25):
        self.config = None
        self.fastq_filenames = None
        self.fastq_sizes = None
        self.fastq_readlength = None
        self.fastq_records_approx = None
        self.spacing = spacing
        self.hits = None
        self.hitseqs = None
        self.stats = None
        self.scantime = 0
        self.coverages = None
        self.results = None
        self.testsuite_versions = {}

    def coverage_at(self, i):                                           # :284-294
        n = len(self.coverages)
        return list(self.coverages.values())[i - n if i >= n else i]

    def get_indexes(self, name):                                        # :296-308
        idx = list(self.coverages.keys()).index(str(name))
        return [idx, idx + len(self.coverages)]

    def __len__(self):
        return len(self.coverages)

    def __getitem__(self, thing):                                       # :310-323: int, Test, or name
        if type(thing) == int:
            return self.coverage_at(thing)
        if hasattr(thing, 'template'):
            return self.coverages[str(thing.template)]
        return self.coverages[str(thing)]

    def scan(self, fastq, templates, do_reverse=True, engine=None, fused=True):        # :325-364
        import time
        if engine is None:
            from . import engine as _e
            engine = _e._engine()
        if templates and all(hasattr(v, 'tests') for v in templates.values()):
            # testsuites instead of bare templates (the reference's signature): load_coverages, :260-282
            from . import genes
            self.testsuites = templates
            self.testsuite_versions = dict((name, ts.version) for name, ts in templates.items())
            templates = genes.template_sequences(templates, self.spacing)
        self.fastq = fastq
        self.fastq_filenames = fastq.filenames()
        self.fastq_sizes = fastq.filesizes()
        self.fastq_readlength = fastq.readlength
        self.fastq_records_approx = fastq.records_approx
        self.coverages = OrderedDict((name, Coverage(seq)) for name, seq in templates.items())
        self.config = engine.get_config()
        seqs = [c.plus_seq.bases for c in self.coverages.values()]
        if do_reverse:
            seqs += [c.minus_seq.bases for c in self.coverages.values()]
        t0 = time.time()
        fused = fused and do_reverse
        ret = engine.findseqs(self.fastq.filenames(), seqs, nplus=len(self.coverages) if fused else 0)
        self.stats = ret['stats']
        self.hits = ret['hits']
        self.hitseqs = ret['hitseqs']
        self.scantime = time.time() - t0
        if fused:
            plus = [c.plus_seq for c in self.coverages.values()]
            for name, c in zip(self.coverages, coverages_from_fused(plus, ret.coverage, ret.mutcount)):
                self.coverages[name] = c
        else:
            self.update_coverages()

    def update_testsuites(self):                                        # :384-395
        self.results = {}
        for name, testsuite in (getattr(self, 'testsuites', None) or {}).items():
            try:
                self.results[name] = testsuite.analyse(self)
            except Exception as e:
                self.results[name] = 'ERROR : ' + str(e)

    def update_coverages(self):                                         # :367-381
        assert self.hits is not None, 'cannot update coverages without .hits'
        assert self.hitseqs is not None, 'cannot update coverages without .hitseqs'
        for hit, hitseq in zip(self.hits, self.hitseqs):
            hit = Hit(*hit)
            self.coverage_at(hit.seq_nr).apply_hit(hit, hitseq, hit.seq_nr < len(self.coverages))

    def encode(self, hits=False):                                       # :398-435
        import time
        more = {}
        if hits:
            more['hits'] = [list(h) for h in self.hits]
            more['hitseqs'] = list(self.hitseqs)
        return dict(
            analyses=self.results,
            info={
                'format': 'kvarq',
                'fastq': self.fastq_filenames,
                'size': self.fastq_sizes,
                'readlength': self.fastq_readlength,
                'records_approx': self.fastq_records_approx,
                'scantime': self.scantime,
                'when': time.asctime(time.localtime()),
                'version': VERSION,
                'config': self.config,
                'spacing': self.spacing,
                'testsuites': dict(self.testsuite_versions),
            },
            stats=self.stats,
            coverages=[(name, c.serialize()) for name, c in self.coverages.items()],
            **more)

    def decode(self, templates, data):                                  # :438-524, current format only
        self.config = data['info']['config']
        self.fastq_filenames = data['info']['fastq']
        self.fastq_sizes = data['info']['size']
        self.fastq_readlength = data['info'].get('readlength', -1)
        self.fastq_records_approx = data['info'].get('records_approx', -1)
        self.stats = data['stats']
        self.scantime = data['info'].get('scantime', -1)
        self.spacing = data['info'].get('spacing', self.spacing)
        self.hits = [Hit(*h) for h in data['hits']] if 'hits' in data else None
        self.hitseqs = data.get('hitseqs')
        self.results = data.get('analyses')
        stored = dict(data['coverages'])
        self.coverages = OrderedDict()
        for name, seq in templates.items():
            c = Coverage(seq)
            if name in stored:
                c.deserialize(stored[name])
                if len(c.coverage) != len(seq):
                    raise ValueError('coverage "%s" has %d positions, template has %d'
                                     % (name, len(c.coverage), len(seq)))
            self.coverages[name] = c
