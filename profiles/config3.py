#!/usr/bin/env python3
"""BASELINE config 3 (gz host streaming path): paired-end 2 x 250 bp MiSeq-profile reads, gzip -6,
scanned end to end from the two .fastq.gz files (host inflate -> pinned batches -> H2D -> kernels).
KVQ_C3_PAIRS pairs (default 300 000 = 2 x 0.17 GB FastQ).  Prints one line per input kind:
the same reads as plain files, as single-member gz, as multi-member gz (one member per 16 MiB).
The gz lines are bound by the host's inflate (zlib, one stream per file) -- say so next to them.
usage: config3.py"""
import gzip, os, sys, time, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvarq_b200 import synth
from kvarq_b200.engine import Engine

npairs = int(os.environ.get('KVQ_C3_PAIRS', 300000))
g = synth.genome()
ts = synth.mtbc_templates(g, seed=1)
iso = synth.isolate(g, ts, seed=7)
tmp = tempfile.mkdtemp(dir='/dev/shm' if os.path.isdir('/dev/shm') else None)
kinds = {'plain': [], 'gz': [], 'gzm': []}
nbases = 0
for k in (1, 2):
    fq = synth.fastq_reads(iso, npairs, readlen=250, seed=3 + k, qmean0=37.0 - 3 * (k - 1), qdrop=20.0, mate=k)
    nbases += synth.count_bases(fq)
    raw = fq.tobytes()
    fn = os.path.join(tmp, 'c3_%d.fastq' % k)
    open(fn, 'wb').write(raw); kinds['plain'].append(fn)
    t0 = time.time()
    z = gzip.compress(raw, 6)
    open(fn + '.gz', 'wb').write(z); kinds['gz'].append(fn + '.gz')
    fm = os.path.join(tmp, 'c3m_%d.fastq.gz' % k)
    with open(fm, 'wb') as f:
        at = 0
        while at < len(raw):
            cut = raw.find(b'\n@SYN', at + (16 << 20))
            cut = len(raw) if cut < 0 else cut + 1
            f.write(gzip.compress(raw[at:cut], 6)); at = cut
    kinds['gzm'].append(fm)
    print('mate %d: %.1f MB FastQ -> %.1f MB gz (%.1f s to compress)' % (k, len(raw) / 1e6, len(z) / 1e6, time.time() - t0), flush=True)

e = Engine(0)
e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.', nthreads=os.cpu_count() or 8)
seqs = ts.seqs()
ref = None
for kind in ('plain', 'gz', 'gzm'):
    e.findseqs(kinds[kind], seqs, nplus=ts.nplus)                  # warm-up (page cache, tables)
    t = []
    for _ in range(3):
        t0 = time.perf_counter()
        r = e.findseqs(kinds[kind], seqs, nplus=ts.nplus)
        t.append(time.perf_counter() - t0)
    key = (len(r['hits']), r['stats']['records_parsed'])
    ref = ref or key
    assert key == ref, (kind, key, ref)
    dt = min(t)
    print('CONFIG3 %-5s e2e %.3f s  %.2f Gbp/s  inflated %.2f GB/s  hits %d records %d kernel_ms %.2f' %
          (kind, dt, nbases / dt / 1e9, sum(os.path.getsize(f) for f in kinds['plain']) / dt / 1e9,
           key[0], key[1], r.timing['kernel_ms']), flush=True)
for fs in kinds.values():
    for f in fs:
        os.remove(f)
os.rmdir(tmp)
