#!/usr/bin/env python3
"""Host-side breakdown of one resident step (KVQ_TRACE=1): usage: trace_step.py"""
import os, sys
os.environ['KVQ_TRACE'] = '1'
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvarq_b200 import synth
from kvarq_b200.engine import Engine
g = synth.genome(); ts = synth.mtbc_templates(g, seed=1)
fq = synth.fastq_reads(synth.isolate(g, ts, seed=7), 3000000, readlen=150, seed=2)
e = Engine(0); e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
st = e.stage([fq])
for _ in range(8):
    r = e.scan_staged(st, ts.seqs(), nplus=ts.nplus, want_hits=True, convert='timing')
