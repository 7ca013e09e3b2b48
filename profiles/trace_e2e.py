#!/usr/bin/env python3
"""Per-batch host timeline of the end-to-end path (KVQ_TRACE=1): usage: trace_e2e.py"""
import os, sys, time
os.environ['KVQ_TRACE'] = '1'
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvarq_b200 import synth
from kvarq_b200.engine import Engine
g = synth.genome(); ts = synth.mtbc_templates(g, seed=1)
fq = synth.fastq_reads(synth.isolate(g, ts, seed=7), 3000000, readlen=150, seed=2)
e = Engine(0); e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
pin = e.pinned_copy(fq)
for i in range(3):
    t = time.perf_counter()
    q = e.findseqs_mem([pin], ts.seqs(), nplus=ts.nplus, want_hits=False)
    sys.stderr.write('[e2e] call %d: %.2f ms\n' % (i, 1e3 * (time.perf_counter() - t)))
