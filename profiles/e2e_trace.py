#!/usr/bin/env python3
"""Timeline of the file path (development tool): engine.findseqs(<config-2 FastQ in /dev/shm>) with KVQ_TRACE=1 --
per batch what the producer (preload, cut) and the consumer (wait, launch, harvest) spend.
usage: KVQ_TRACE=1 e2e_trace.py [reads]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvarq_b200 import synth
from kvarq_b200.engine import Engine
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000000
g = synth.genome()
ts = synth.mtbc_templates(g, seed=1)
fq = synth.fastq_reads(synth.isolate(g, ts, seed=7), n, readlen=150, seed=2)
fn = '/dev/shm/kvq_e2e_trace.fastq'
open(fn, 'wb').write(memoryview(fq))
e = Engine(0)
e.config(nthreads=os.cpu_count() or 8, maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
os.environ.pop('KVQ_TRACE_ON', None)
for rep in range(3):
    sys.stderr.write('[e2e_trace] ---- call %d\n' % rep); sys.stderr.flush()
    t = time.perf_counter()
    r = e.findseqs(fn, ts.seqs(), nplus=ts.nplus, convert=False)
    dt = time.perf_counter() - t
    print('E2E_TRACE call %d: %.2f ms, %.1f GB/s, %d hits' % (rep, 1e3 * dt, fq.nbytes / dt / 1e9, r.nhits), flush=True)
os.remove(fn)
