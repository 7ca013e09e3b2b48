#!/usr/bin/env python3
"""BASELINE config 4 (template scaling): kernel time of the resident scan as the template set grows
from MTBC (264 seqs) by synthetic 51-bp SNP panels of 1 024 / 4 096 / 16 384 positions (seed 4).
Reads: the config-2 recipe, KVQ_C4_READS of them (default 1 000 000 = 0.33 GB, larger than L2).
usage: config4.py [panel sizes ...]      prints one line per panel size."""
import os, sys, time, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvarq_b200 import synth
from kvarq_b200.engine import Engine

sizes = [int(x) for x in sys.argv[1:]] or [0, 1024, 4096, 16384]
nreads = int(os.environ.get('KVQ_C4_READS', 1000000))
g = synth.genome()
base = synth.mtbc_templates(g, seed=1)
fq = synth.fastq_reads(synth.isolate(g, base, seed=7), nreads, readlen=150, seed=2)
nbases = synth.count_bases(fq)
e = Engine(0)
e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
st = e.stage([fq])
for n in sizes:
    plus = list(base.plus)
    if n:
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
    print('CONFIG4 panel %6d seqs %6d kernel_ms %9.3f Gbp/s %8.2f GB/s %8.1f hits %8d candidates %11d wall_s %.1f'
          % (n, len(seqs), ms, nbases / ms / 1e6, fq.nbytes / ms / 1e6, r.nhits, r.timing['candidates'],
             time.time() - t0), flush=True)
