#!/usr/bin/env python
"""Generates tests/golden/{ref_inputs.tar.gz, golden.json} -- run in the BUILD container only.

Inputs : the reference's own FastQ fixtures (/root/reference/tests/fastqs/*) and the 43 spoligo
         spacers (literal 25-mers in /root/reference/testsuites/MTBC/spoligo.py:75-119), bundled
         unchanged into ref_inputs.tar.gz because /root/reference does not exist on the GPU box.
Outputs: for every case (the configurations of the reference's tests/test_engine.py and
         tests/test_analyser.py plus default-config spoligo scans of the bundled real reads) the
         full result of the REFERENCE engine -- csrc/workhorse.c compiled unmodified
         (oracle/_ref/libkvarq_ref.so, nthreads=1) -- hits, hitseqs and stats.
"""
import io
import json
import os
import re
import sys
import tarfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import checker  # noqa: E402

REF = '/root/reference'
FQ = os.path.join(REF, 'tests', 'fastqs')

ENGINE_SEQS = ["CCC", "TTTT", "TATATATA", "TGTAG", "ATATT",
               "GAGCATGTGGAGCAACTTGTGGGAGCGCCGGGCAACGCCCTGTCTCTTAT",
               "...NACTTCCTCTCTACTGGTGTCGGCGGTGAAAGAGCTTACGTACTCTTCGAT..."]
MAXERR_SEQS = ["CAGCATGTGGAGCAACTTGTGGGAGCGCCGGGCAACGCCCTGTCTCTTAT",
               "CTGCATGTGGAGCAACTTGTGGGAGCGCCGGGCAACGCCCTGTCTCTTAT",
               "CTCCATGTGGAGCAACTTGTGGGAGCGCCGGGCAACGCCCTGTCTCTTAT"]
MINOVL_SEQS = ["TCGATGCGATCTGTCAAGTCGGTGGCGGTA...", "TCGATGCGATCTG.CAAGTCGGTGGCGGTA...",
               "...NTGAACGTATCGCCTCGAGGGACTT", "...NTGAACGTATCG.CTCGAGGGACTT"]


def revcomp(s):
    return s.translate(str.maketrans('ACGTN', 'TGCAN'))[::-1]


def spacers():
    src = open(os.path.join(REF, 'testsuites', 'MTBC', 'spoligo.py')).read()
    sp = re.findall(r"StaticTemplate\('([ACGT]{25})'\)", src)
    assert len(sp) == 43
    return sp


def main():
    checker.build()
    sp = spacers()
    spol = sp + [revcomp(s) for s in sp]            # kvarq/analyse.py:352-354
    cases = []

    def case(name, files, seqs, **cfg):
        cases.append(dict(name=name, files=files, seqs=seqs, cfg=cfg))

    base = dict(maxerrors=0, minoverlap=1000, minreadlength=3, Amin='!')
    case('findseqs', ['test_engine.fastq'], ENGINE_SEQS, **base)                    # test_engine.py:137
    case('findseqs_gz', ['test_engine.fastq.gz'], ENGINE_SEQS, **base)              # :178
    case('paired', ['test_engine_1.fastq', 'test_engine_2.fastq'], ENGINE_SEQS, **base)   # :184
    case('paired_gz', ['test_engine_1.fastq.gz', 'test_engine_2.fastq.gz'], ENGINE_SEQS, **base)
    for e in range(4):                                                              # :208
        case('maxerror_%d' % e, ['test_engine.fastq'], MAXERR_SEQS,
             maxerrors=e, minoverlap=25, minreadlength=25, Amin='!')
    case('minoverlap_30', ['test_engine.fastq'], MINOVL_SEQS, maxerrors=0, minoverlap=30, minreadlength=25, Amin='!')
    case('minoverlap_25', ['test_engine.fastq'], MINOVL_SEQS, maxerrors=0, minoverlap=25, minreadlength=25, Amin='!')
    case('minoverlap_25_e1', ['test_engine.fastq'], MINOVL_SEQS, maxerrors=1, minoverlap=25, minreadlength=25, Amin='!')
    case('Amin_H', ['test_engine.fastq'], ["GGAG", "CCGAC"], maxerrors=0, minoverlap=25, minreadlength=4, Amin='H')   # :257
    case('Amin_G', ['test_engine.fastq'], ["GGAG", "CCGAC"], maxerrors=0, minoverlap=25, minreadlength=4, Amin='G')
    # tests/test_analyser.py:57-64 (spoligo part)
    case('analyser_spoligo', ['test_analyser.fastq'], spol, maxerrors=1, minoverlap=10, minreadlength=25, Amin='!')
    # product defaults (kvarq/config.py:2-10; Q13 Sanger -> Amin '.') on the bundled real reads
    dflt = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
    for f in ['L3_N1014_hits_5k.fastq', 'N0116_1_hits_1k.fastq', 'test_analyser.fastq']:
        case('spoligo_default_' + f.split('.')[0], [f], spol, **dflt)
    case('spoligo_default_pair', ['L3_N1014_hits_500_1.fastq', 'L3_N1014_hits_500_2.fastq'], spol, **dflt)
    case('spoligo_e3_5k', ['L3_N1014_hits_5k.fastq'], spol, maxerrors=3, minoverlap=20, minreadlength=20, Amin='.')

    out = []
    for c in cases:
        paths = [os.path.join(FQ, f) for f in c['files']]
        r = checker.findseqs(paths, c['seqs'], nthreads=1, impl='ref', **c['cfg'])
        o = checker.findseqs(paths, c['seqs'], nthreads=1, impl='oracle', **c['cfg'])
        assert r == o, 'restatement differs from the reference on ' + c['name']
        out.append(dict(name=c['name'], files=c['files'], seqs=c['seqs'], cfg=c['cfg'],
                        hits=[list(h) for h in r['hits']],
                        hitseqs=[h.decode('latin-1') for h in r['hitseqs']],
                        stats={k: (list(v) if isinstance(v, tuple) else v) for k, v in r['stats'].items()}))
        print('%-32s %5d hits  %6d records' % (c['name'], len(r['hits']), r['stats']['records_parsed']))

    with tarfile.open(os.path.join(HERE, 'ref_inputs.tar.gz'), 'w:gz') as tar:
        for f in sorted(os.listdir(FQ)):
            tar.add(os.path.join(FQ, f), arcname=f)
    json.dump(dict(generator='tests/golden/make_golden.py', reference_commit='d693f56',
                   spacers=sp, cases=out), open(os.path.join(HERE, 'golden.json'), 'w'), indent=0)


if __name__ == '__main__':
    main()
