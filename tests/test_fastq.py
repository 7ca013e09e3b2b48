"""Pre-flight: the cases of the reference's tests/test_fastq.py:45-141 against
the restated `kvarq_b200.fastq.Fastq`, plain and gzipped, plus the record helpers on Hit.file_pos."""
import gzip
import logging
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from kvarq_b200.fastq import Fastq, FastqFileFormatException
from kvarq_b200.analyse import Hit


def _write(tmp_path, content, gz, variant=None, name='t'):
    fn = str(tmp_path / (name + '.fastq' + ('.gz' if gz else '')))
    with (gzip.GzipFile(fn, 'w') if gz else open(fn, 'wb')) as f:
        f.write(content.encode('latin-1'))
    return Fastq(fn, variant=variant, quiet=True)


def _quality(tmp_path, quality, gz, variant=None):
    return _write(tmp_path, '@IDENTIFIER\n' + 'A' * len(quality) + '\n+\n' + quality + '\n', gz, variant)


@pytest.mark.parametrize('gz', [False, True])
def test_fastq_variant(tmp_path, gz, caplog):
    """tests/test_fastq.py:45-80"""
    fq = _quality(tmp_path, '!"#$%&\'()*+,-./0123456789:;<=>?@ABCDEFGHIJ', gz)
    assert fq.dQ == 0 and set(fq.variants) == set(['Illumina 1.8+', 'Sanger']) and fq.Azero == '!'
    fq = _quality(tmp_path, ';<=>?@ABCDEFGHIJKLMNOPQRSTUVWXYZ[\\]^_`abcdefgh', gz)
    assert fq.dQ == 31 and fq.variants == ['Solexa'] and fq.Azero == '@'
    fq = _quality(tmp_path, 'OPQRSTUVWXYZ[\\]^_`abcdefgh', gz)
    assert fq.dQ == 31 and fq.variants == ['Solexa', 'Illumina 1.3+', 'Illumina 1.5+']
    with pytest.raises(FastqFileFormatException):
        _quality(tmp_path, ';<=>?@ABCDEFGHI;<=>?@ABCDEFGHI', gz)        # ambiguous dQ
    _quality(tmp_path, ';<=>?@ABCDEFGHI;<=>?@ABCDEFGHI', gz, variant='Sanger')
    fq = _quality(tmp_path, ';<=>?@ABCDEFGHI;<=>?@ABCDEFGHI', gz, variant='Solexa')
    assert fq.dQ == 31 and fq.Q2A(13) == 'M' and fq.A2Q('M') == 13
    with caplog.at_level(logging.WARNING, logger='kvarq_b200'):
        _quality(tmp_path, ';<=>?@ABCDEFGHI;<=>?@ABCDEFGHI', gz, variant='Illumina 1.3+')
    assert 'seems not to be compatible' in caplog.text
    with pytest.raises(FastqFileFormatException):
        _quality(tmp_path, 'IIII', gz, variant='no such vendor')


BASES = 'ACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGT'
BAD = [
    'IDENTIFIER\n' + BASES + '\n+\n' + '#' * 44 + '\n',                         # no '@'
    '@IDENTIFIER\n' + BASES.replace('ACGTACGTACGTACGTACGTAC', 'ACGTACGTACGTACGTACGTAX') + '\n+\n' + '#' * 44 + '\n',
    '@IDENTIFIER\n' + BASES + '\n+\n' + '#' * 45 + '\n',                        # one score too many
    '@IDENTIFIER\n' + BASES + '\n+text\n' + '#' * 44 + '\n',                    # '+' line not '+' / '+ident'
    '@IDENTIFIER\n' + BASES + '\n+\n' + '#' * 44 + '\n\n@IDENTIFIER\n',         # text after an empty line
]


@pytest.mark.parametrize('gz', [False, True])
@pytest.mark.parametrize('case', range(len(BAD)))
def test_fastq_format(tmp_path, gz, case):
    """tests/test_fastq.py:83-135"""
    with pytest.raises(FastqFileFormatException):
        _write(tmp_path, BAD[case], gz)


def test_accepted_shapes(tmp_path):
    """'+IDENTIFIER' separator, CRLF, one trailing '!' more than bases (kvarq/fastq.py:208-216);
    suffix rule (:81-93); empty file (:105)"""
    fq = _write(tmp_path, '@ID\r\n' + 'A' * 30 + '\r\n+ID\r\n' + 'I' * 30 + '!\r\n', False)
    assert fq.readlength == 30 and fq.records_approx == 1
    fn = str(tmp_path / 'reads.fq')
    open(fn, 'w').write('@ID\nA\n+\nI\n')
    with pytest.raises(FastqFileFormatException):
        Fastq(fn, quiet=True)
    fn = str(tmp_path / 'empty.fastq')
    open(fn, 'w').close()
    with pytest.raises(FastqFileFormatException):
        Fastq(fn, quiet=True)


def test_pairing_and_sizes(tmp_path):
    """kvarq/fastq.py:95-103: `x_1.fastq` pulls in `x_2.fastq` when paired=True and it exists"""
    rec = '@r\n' + 'ACGT' * 10 + '\n+\n' + '#' + 'I' * 39 + '\n'
    f1, f2 = str(tmp_path / 's_1.fastq'), str(tmp_path / 's_2.fastq')
    open(f1, 'w').write(rec * 3)
    open(f2, 'w').write(rec * 2)
    fq = Fastq(f1, paired=True, quiet=True)
    assert fq.filenames() == [f1, f2] and fq.filesizes() == [3 * len(rec), 2 * len(rec)]
    assert fq.records_approx == 6
    assert Fastq(f1, paired=False, quiet=True).filenames() == [f1]
    assert Fastq(f2, paired=True, quiet=True).filenames() == [f2]


def test_cutoff_and_record_helpers(tmp_path):
    """cutoff (:301-314) closes a run only on a worse score; readhit / readrecordat work on the
    engine's file_pos (= first base of the trimmed read)"""
    recs = ['@r%d\n%s\n+\n%s\n' % (i, 'ACGT' * 5 + 'TTTTT' * i, '5' + 'I' * (19 + 5 * i)) for i in range(1, 6)]
    text = ''.join(recs)
    fn = str(tmp_path / 'h.fastq')
    open(fn, 'w').write(text)
    fq = Fastq(fn, quiet=True)
    assert fq.cutoff('IIII##IIIIII#', '.') == (6, 6)
    assert fq.cutoff('IIIIII', '.') == (0, -1)          # never closed: not counted (as written there)
    assert fq.lengths('.', n=5, points=1) == []
    pos = text.index(recs[2]) + len('@r3\n')
    hit = Hit(seq_nr=0, file_pos=pos, seq_pos=-4, length=10, readlength=35)
    assert fq.readhit(hit) == recs[2].split('\n')[1][4:14]
    assert fq.readhit(hit._replace(seq_pos=3)) == recs[2].split('\n')[1][:10]
    assert fq.readrecordat(hit) == recs[2]
    assert fq.p2Q(.05) == 13 and abs(fq.Q2p(13) - 0.0501) < 1e-3
