#!/usr/bin/env python3
"""Kernel-variant sweep (development tool): times kvq_scan_staged for several builds of the library
on the bench workload.  usage: sweep.py tag1 tag2 ...   (libs: /root/repo/build_variants/lib_<tag>.so)"""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CHILD = r'''
import os, sys, numpy as np, pickle
sys.path.insert(0, %r)
from kvarq_b200.engine import Engine
fq = np.load('/dev/shm/kvq_fq.npy', mmap_mode='r'); seqs, nplus = pickle.load(open('/dev/shm/kvq_ts.pkl', 'rb'))
e = Engine(0); e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
st = e.stage([np.ascontiguousarray(fq)])
for _ in range(3): r = e.scan_staged(st, seqs, nplus=nplus, want_hits=True, convert='timing')
k = []
for _ in range(7):
    r = e.scan_staged(st, seqs, nplus=nplus, want_hits=True, convert='timing'); k.append(r.timing['kernel_ms'])
print('RESULT %%s kernel_ms min %%.4f med %%.4f hits %%d cands %%d' %% (os.environ.get('KVQ_B200_LIB', 'default').split('/')[-1], min(k), sorted(k)[3], r.nhits, r.timing['candidates']))
''' % ROOT
if not os.path.exists('/dev/shm/kvq_fq.npy'):
    import numpy as np, pickle
    from kvarq_b200 import synth
    g = synth.genome(); ts = synth.mtbc_templates(g, seed=1)
    fq = synth.fastq_reads(synth.isolate(g, ts, seed=7), int(os.environ.get('KVQ_BENCH_READS', 3000000)), readlen=150, seed=2)
    np.save('/dev/shm/kvq_fq.npy', fq); pickle.dump((ts.seqs(), ts.nplus), open('/dev/shm/kvq_ts.pkl', 'wb'))
for tag in sys.argv[1:]:
    env = dict(os.environ)
    if tag != 'default': env['KVQ_B200_LIB'] = os.path.join(ROOT, 'build_variants', 'lib_%s.so' % tag)
    out = subprocess.run([sys.executable, '-c', CHILD], env=env, capture_output=True, text=True)
    print([l for l in out.stdout.split('\n') if l.startswith('RESULT')] or out.stderr[-800:], flush=True)
