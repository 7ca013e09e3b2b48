#!/usr/bin/env python3
"""BASELINE config 4 (template scaling): kernel time of the resident scan as the template set grows
from MTBC (264 seqs) by the REAL barcode sets -- Coll-2014 (62 SNPs, testsuites/MTBC-SNP-barcodes/coll14.tsv, on
the synthetic ancestor) and CanSNPer francisella + coxiella (75 SNP rows at their positions on the shipped
genomes, testsuites/CanSNPer/*_snp.txt; tables + genome neighbourhoods in tests/golden/testsuite_data.json):
panel "real" -- and by synthetic 51-bp SNP panels of 1 024 / 4 096 / 16 384 positions (seed 4) on top.
Reads: the config-2 recipe, KVQ_C4_READS of them (default 1 000 000 = 0.33 GB, larger than L2).
usage: config4.py [panel sizes ...]      prints one line per panel size."""
import os, sys, time, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvarq_b200 import synth
from kvarq_b200.engine import Engine

sizes = [x if x == 'real' else int(x) for x in sys.argv[1:]] or [0, 'real', 1024, 4096, 16384]
nreads = int(os.environ.get('KVQ_C4_READS', 1000000))
g = synth.genome()
base = synth.mtbc_templates(g, seed=1)
fq = synth.fastq_reads(synth.isolate(g, base, seed=7), nreads, readlen=150, seed=2)
nbases = synth.count_bases(fq)
e = Engine(0)
e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
st = e.stage([fq])


def real_barcodes():
    """plus-strand templates (51 bp, mutant base in the middle: SNP.seq, kvarq/genes.py:498-501) of coll14 and the
    two CanSNPer organisms whose genomes the reference ships"""
    import json
    from kvarq_b200 import genes, testsuites
    data = json.load(open(os.path.join(ROOT, 'tests', 'golden', 'testsuite_data.json')))
    G = genes.Genome.from_bases(g, 'synthetic ancestor')
    suites = {'coll14': testsuites.barcode(G, data['coll14'], force=True)}
    for org in ('francisella_tularensis', 'coxiella_burnetii'):
        d = data['cansnper'][org]
        gs = dict((s, genes.SparseGenome(w['windows'], s, size=w['size'])) for s, w in d['windows'].items())
        suites[org] = testsuites.snp_tree(gs, d['snps'], d['tree'])
    return [s.bases.encode() for s in genes.template_sequences(suites, 25).values()]


real = None
for n in sizes:
    plus = list(base.plus)
    if n == 'real' or (n and 'real' in sizes):
        real = real or real_barcodes()
        plus += real
    if n and n != 'real':
        plus += list(synth.snp_panel(g, n, seed=4).plus)
    seqs = plus + [synth.revcomp(s) for s in plus]
    t0 = time.time()
    for _ in range(2):
        r = e.scan_staged(st, seqs, nplus=len(plus), want_hits=True, convert='timing')
    k = []
    for _ in range(5):
        r = e.scan_staged(st, seqs, nplus=len(plus), want_hits=True, convert='timing')
        k.append(r.timing['kernel_ms'])
    ms = sorted(k)[2]
    print('CONFIG4 panel %6s seqs %6d kernel_ms %9.3f Gbp/s %8.2f GB/s %8.1f hits %8d candidates %11d wall_s %.1f'
          % (n, len(seqs), ms, nbases / ms / 1e6, fq.nbytes / ms / 1e6, r.nhits, r.timing['candidates'],
             time.time() - t0), flush=True)
