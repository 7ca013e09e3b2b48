"""Host-side pre-flight of a scan: Python-3 restatement of the reference's
``kvarq.fastq`` -- `FastqFileFormatException` (the type the engine raises, imported by
csrc/workhorse.c:1598-1600) and `Fastq` (kvarq/fastq.py:43-387): `.fastq` / `.fastq.gz` suffix
rule, `_1` / `_2` pairing, format check of sampled records, vendor-variant sniffing -> `dQ`,
`Azero`, `Q2A()` (what `kvarq scan` feeds into `engine.config(Amin=..., Azero=...)`,
kvarq/cli.py:73-86), and the record helpers that work on `Hit.file_pos`.

Files are read as bytes and decoded latin-1, so every method sees the same characters the Python-2
original did.  Python-2 integer divisions (`size*point/points`, `(points-1-point)*n/points`,
`records_approx`) are kept as `//`.
"""
import collections
import gzip
import logging
import math
import os.path

lo = logging.getLogger('kvarq_b200')


class FastqFileFormatException(Exception):
    """Raised for records not starting with '@' / third lines not starting with '+'
    (kvarq/fastq.py:40, raised from csrc/workhorse.c:1037-1048)."""


class _Text(object):
    """binary file seen as latin-1 text (Python-2 `str` semantics for readline/read/seek/tell)"""

    def __init__(self, fd):
        self.fd = fd

    def readline(self):
        return self.fd.readline().decode('latin-1')

    def read(self, n=-1):
        return self.fd.read(n).decode('latin-1')

    def seek(self, pos):
        return self.fd.seek(pos)

    def tell(self):
        return self.fd.tell()

    def close(self):
        self.fd.close()


class Fastq(object):
    """kvarq/fastq.py:43-387"""

    ASCII = '!"#$%&\'()*+,-./0123456789:;<=>?@ABCDEFGHIJKLMNOPQRSTUVWXYZ' + \
            '[\\]^_`abcdefghijklmnopqrstuvwxyz{|}~'

    VendorProperties = collections.namedtuple('VendorProperties', ['Qrange', 'dQ'])

    # :49-55 (a dict there; the order below reproduces the lists tests/test_fastq.py:52-58 expect)
    vendor_variants = collections.OrderedDict((
        ('Sanger', VendorProperties(range(0, 50), 0)),
        ('Solexa', VendorProperties(range(-5, 41), 31)),
        ('Illumina 1.3+', VendorProperties(range(0, 41), 31)),
        ('Illumina 1.5+', VendorProperties(range(3, 42), 31)),
        ('Illumina 1.8+', VendorProperties(range(0, 62), 0)),
    ))

    def __init__(self, fname, variant=None, fd=None, paired=False, quiet=False):      # :57-160
        self.fname = fname
        self.fd = _Text(fd) if fd else None
        if self.fname.endswith('.fastq.gz'):
            self.gz = True
            if not self.fd:
                self.fd = _Text(gzip.GzipFile(self.fname, 'rb'))
        elif self.fname.endswith('.fastq'):
            self.gz = False
            if not self.fd:
                self.fd = _Text(open(self.fname, 'rb'))
        else:
            raise FastqFileFormatException('fastq file must have extension ".fastq" or ".fastq.gz"')

        self.fname2 = None
        if paired:
            base = fname[:fname.rindex('.fastq')]
            if base[-2:] == '_1':
                fname2 = base[:-2] + '_2' + fname[fname.rindex('.fastq'):]
                if os.path.exists(fname2):
                    lo.info('including paired file "%s"' % fname2)
                    self.fname2 = fname2

        if sum(self.filesizes()) == 0:
            raise FastqFileFormatException('cannot scan empty file')

        min_pos, max_pos = self.min_max_score_check_file()
        lo.debug('min_pos=%d max_pos=%d' % (min_pos, max_pos))

        if variant and variant not in self.vendor_variants:
            raise FastqFileFormatException('unknown vendor variant "%s"' % variant)

        variants, dQs = [], []
        for name, vv in self.vendor_variants.items():
            if (min_pos - vv.dQ) in vv.Qrange and (max_pos - vv.dQ) in vv.Qrange:
                dQs.append(vv.dQ)
                variants.append(name)

        if variant is None:
            if not variants:
                raise FastqFileFormatException('could not find any suitable fastq vendor variant')
            if len(set(dQs)) > 1:
                raise FastqFileFormatException(
                    'cannot determine dQ with guessed vendor variants "%s"' % str(variants))
            self.variants = variants
            self.dQ = dQs[0]
        else:
            if variant not in variants:
                lo.warning('specified vendor variant "%s" seems not to be compatible with file' % variant)
            self.variants = [variant]
            self.dQ = self.vendor_variants[variant].dQ

        self.Azero = self.ASCII[self.dQ]

        self.fd.seek(0)
        lines = [self.fd.readline() for i in range(4)]
        self.readlength = len(lines[1].strip('\r\n'))
        if self.gz:
            self.records_approx = None
        else:
            self.records_approx = os.path.getsize(self.fname) // len(''.join(lines))
            if self.fname2 is not None:
                self.records_approx *= 2

        if not quiet:
            if self.gz:
                lo.info('gzipped fastq : readlength=? records_approx=? dQ=%d variants=%s' % (
                    self.dQ, str(self.variants)))
            else:
                lo.info('fastq : readlength=%d records_approx=%d dQ=%d variants=%s' % (
                    self.readlength, self.records_approx, self.dQ, str(self.variants)))

    def filesizes(self):                                                              # :163-166
        return [os.path.getsize(fname) for fname in self.filenames()]

    def filenames(self):                                                              # :168-173
        if self.fname2 is not None:
            return [self.fname, self.fname2]
        return [self.fname]

    def min_max_score_check_file(self, n=1000, points=10):                           # :175-243
        """format check of n records sampled at `points` places (start only for .gz);
        returns min / max quality character as indices into ASCII"""
        ret_min = +999
        ret_max = -999
        self.fd.seek(0)
        if self.gz:
            lo.debug('gzipped fastq : scan %d points at start only' % n)
        identifier = None
        for point in range(points):
            if not self.gz and point > 0:
                self.fd.seek(os.path.getsize(self.fname) * point // points)
                self.seekback()
            while n > (points - 1 - point) * n // points:
                identifier = self.fd.readline().rstrip('\n\r')
                if not identifier:
                    break
                if not identifier[0] == '@':
                    raise FastqFileFormatException('identifier (1st line of record) must begin with "@"')
                bases = self.fd.readline().rstrip('\n\r')
                if not set(bases).issubset(set('AGCTN')):
                    raise FastqFileFormatException('bases (2nd line of record) must contain only AGCTN')
                plus = self.fd.readline().rstrip('\n\r')
                if not (plus == '+' or (plus[0] == '+' and plus[1:] == identifier[1:])):
                    raise FastqFileFormatException(
                        'separator (3rd line of record) must be == "+" or "+(ident)"')
                phredstr = self.fd.readline().rstrip('\n\r')
                if not (len(bases) == len(phredstr) or (
                        len(bases) == len(phredstr) - 1 and phredstr[-1] == '!')):
                    raise FastqFileFormatException(
                        'bases must be ~ same length as phred score (2nd, 4th line)')
                try:
                    idx = [self.ASCII.index(x) for x in phredstr]
                except ValueError:
                    raise FastqFileFormatException(
                        'phred score (4th line of record) must contain only "%s"' % self.ASCII)
                ret_min = min([ret_min] + idx)
                ret_max = max([ret_max] + idx)
                n -= 1
            if not identifier:
                break
        if not identifier:
            while True:
                line = self.fd.readline()
                if not line:
                    break
                if not line.rstrip('\r\n') == '':
                    raise FastqFileFormatException('non-empty line after empty line (fpos=%d' % self.fd.tell())
        return ret_min, ret_max

    def A2Q(self, A):                                                                 # :246-252
        return self.ASCII.index(A) - self.dQ

    def Q2A(self, Q):
        return self.ASCII[Q + self.dQ]

    def Q2p(self, Q):                                                                 # :255-266
        return 10 ** (-.1 * Q)

    def p2Q(self, p):
        return int(-10 * math.log(p) / math.log(10))

    def lengths(self, Amin, n=1000, points=10):                                       # :269-299
        """lengths of the quality-trimmed reads of n sampled records"""
        self.fd.seek(0)
        lengths = []
        for point in range(points):
            if not self.gz and point > 0:
                self.fd.seek(os.path.getsize(self.fname) * point // points)
                self.seekback()
            while n > (points - 1 - point) * n // points:
                ident, seq, plus, scores = [self.fd.readline().strip() for j in range(4)]
                pos, length = self.cutoff(scores, Amin)
                if length >= 0:
                    lengths.append(length)
                n -= 1
        return lengths

    def cutoff(self, scores, Amin):                                                   # :301-314
        """(pos, length) of the first longest run >= Amin that is closed by a worse score
        (as written there: a run reaching the end of `scores` is not counted)"""
        length = -1
        pos_ = pos = 0
        for j, A in enumerate(scores):
            if ord(A) >= ord(Amin):
                if pos < 0:
                    pos = j
            else:
                if pos >= 0 and length < j - pos:
                    length = j - pos
                    pos_ = pos
                pos = -1
        return pos_, length

    def readhit(self, hit):                                                           # :316-324
        if hit.seq_pos < 0:
            self.fd.seek(hit.file_pos - hit.seq_pos)
            return self.fd.read(hit.length)
        self.fd.seek(hit.file_pos)
        return self.fd.read(hit.length)

    def readhits(self, hits):                                                         # :384-386
        return [self.readhit(hit) for hit in hits]

    def lineup(self):                                                                 # :326-337
        """file position -> beginning of the current line (of the previous one if already there)"""
        pos = self.fd.tell()
        c = None
        while c != '\n' and pos > 0:
            pos -= 1
            self.fd.seek(max(0, pos - 1))
            c = self.fd.read(1)
        if pos == 0:
            self.fd.seek(0)

    def seekback(self):                                                               # :339-359
        """file position -> beginning of the current (after the '+' line) or previous record"""
        l = pos = None
        while pos != 0:
            self.lineup()
            l = self.fd.readline()
            self.lineup()
            if l[0] == '+':
                self.lineup()
                l = self.fd.readline()          # a '+' may also start a quality line ...
                self.lineup()
                if l[0] == '+':
                    self.lineup()               # ... then the separator is one line further up
                self.lineup()
                break
            pos = self.fd.tell()

    def readrecord(self):                                                             # :376-381
        ident, seq, plus, scores = [self.fd.readline().strip() for j in range(4)]
        return ident, seq, plus, scores

    def readrecordat(self, hit):                                                      # :383-390
        self.fd.seek(hit.file_pos)
        self.seekback()
        ident, seq, plus, scores = self.readrecord()    # previous record
        ident, seq, plus, scores = self.readrecord()    # the hit's record
        return '\n'.join([ident, seq, plus, scores]) + '\n'

    def dumpat(self, pos, Amin=None):                                                 # :361-374
        self.fd.seek(pos)
        self.seekback()
        ident, seq, plus, scores = self.readrecord()
        print(ident)
        print('-'.join(['%s(%d)' % (base, self.A2Q(A)) for base, A in zip(seq, scores)]))
        if Amin:
            pos, length = self.cutoff(scores, Amin)
            print('Amin=%s -> pos=%d length=%d' % (Amin, pos, length))
