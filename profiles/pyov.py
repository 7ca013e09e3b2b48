import os, sys, time, numpy as np, pickle, cProfile, pstats
sys.path.insert(0, '/root/repo')
from kvarq_b200.engine import Engine
from kvarq_b200 import synth
g = synth.genome(); ts = synth.mtbc_templates(g, seed=1)
fq = synth.fastq_reads(synth.isolate(g, ts, seed=7), 3000000, readlen=150, seed=2)
seqs = ts.seqs()
e = Engine(0); e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
st = e.stage([fq])
for _ in range(3): r = e.scan_staged(st, seqs, nplus=ts.nplus, want_hits=True, convert='timing')
t0 = time.perf_counter(); w = 0
for _ in range(20):
    r = e.scan_staged(st, seqs, nplus=ts.nplus, want_hits=True, convert='timing'); w += r.timing['wall_ms']
dt = (time.perf_counter() - t0) / 20 * 1e3
print('py wall %.3f ms, C wall %.3f ms, kernel %.3f' % (dt, w / 20, r.timing['kernel_ms']))
pr = cProfile.Profile(); pr.enable()
for _ in range(20): r = e.scan_staged(st, seqs, nplus=ts.nplus, want_hits=True, convert='timing')
pr.disable(); pstats.Stats(pr).sort_stats('cumulative').print_stats(8)
