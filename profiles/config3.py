#!/usr/bin/env python3
"""BASELINE config 3 (gz host streaming path): paired-end 2 x 250 bp MiSeq-profile reads, gzip -6,
scanned end to end from the two .fastq.gz files (host inflate -> pinned batches -> H2D -> kernels).
KVQ_C3_PAIRS pairs (default: the stated 1 764 613 = 200x, 2 x 0.92 GB FastQ).  Prints one line per
input kind: the same reads as plain files, as single-member gz (chunk-parallel inflate, gz_parallel.h),
as single-member gz with one zlib stream per file (KVQ_GZ_THREADS=1, the round-1 path) and -- with
KVQ_C3_GZM=1 -- as multi-member gz (one member per 16 MiB) and as BGZF-like gz (one member per 64 KiB); KVQ_C3_SWEEP=12,14 adds lines with that many inflate threads.
Parity at this size: hits / hitseqs / stats of every gz scan == those of the plain-file scan of the same
reads (the plain path is pinned to the reference by tests/), and the first KVQ_C3_ORACLE_PAIRS pairs,
gzipped on their own, are checked against the CPU checker (oracle/, nthreads = cores).
usage: config3.py"""
import gzip, os, sys, time, tempfile
from concurrent.futures import ProcessPoolExecutor
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kvarq_b200 import synth

npairs = int(os.environ.get('KVQ_C3_PAIRS', 1764613))
npre = int(os.environ.get('KVQ_C3_ORACLE_PAIRS', 60000))
CFG = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')


def make_mate(args):
    k, tmp, with_members = args
    g = synth.genome()
    ts = synth.mtbc_templates(g, seed=1)
    iso = synth.isolate(g, ts, seed=7)
    fq = synth.fastq_reads(iso, npairs, readlen=250, seed=3 + k, qmean0=37.0 - 3 * (k - 1), qdrop=20.0, mate=k)
    nb = synth.count_bases(fq)
    raw = fq.tobytes()
    fn = os.path.join(tmp, 'c3_%d.fastq' % k)
    open(fn, 'wb').write(raw)
    t0 = time.time()
    z = gzip.compress(raw, 6)
    open(fn + '.gz', 'wb').write(z)
    out = {'plain': fn, 'gz': fn + '.gz', 'bases': nb, 'raw': len(raw), 'z': len(z), 'tz': time.time() - t0}
    cut = raw.find(b'\n@SYN', npre * 540) + 1 if npre * 600 < len(raw) else len(raw)
    nl = 0; at = 0
    for _ in range(4 * npre):               # exactly npre records
        at = raw.find(b'\n', at) + 1
    pre = os.path.join(tmp, 'c3pre_%d.fastq.gz' % k)
    open(pre, 'wb').write(gzip.compress(raw[:at], 6))
    out['pre'] = pre
    if with_members:
        fm = os.path.join(tmp, 'c3m_%d.fastq.gz' % k)
        with open(fm, 'wb') as f:
            at = 0
            while at < len(raw):
                cut = raw.find(b'\n@SYN', at + (16 << 20))
                cut = len(raw) if cut < 0 else cut + 1
                f.write(gzip.compress(raw[at:cut], 6)); at = cut
        out['gzm'] = fm
        fb = os.path.join(tmp, 'c3b_%d.fastq.gz' % k)          # BGZF-like: one member per 64 KiB (Illumina's converters)
        with open(fb, 'wb') as f:
            for at in range(0, len(raw), 65280):
                f.write(gzip.compress(raw[at:at + 65280], 6))
        out['bgzf'] = fb
    return out


def main():
    from kvarq_b200.engine import Engine
    tmp = tempfile.mkdtemp(dir='/dev/shm' if os.path.isdir('/dev/shm') else None)
    gzm = os.environ.get('KVQ_C3_GZM') == '1'
    with ProcessPoolExecutor(2) as ex:
        mates = list(ex.map(make_mate, [(1, tmp, gzm), (2, tmp, gzm)]))
    for k, m in enumerate(mates):
        print('mate %d: %.1f MB FastQ -> %.1f MB gz (%.1f s to compress)' % (k + 1, m['raw'] / 1e6, m['z'] / 1e6, m['tz']), flush=True)
    nbases = sum(m['bases'] for m in mates)
    raw_total = sum(m['raw'] for m in mates)
    g = synth.genome()
    ts = synth.mtbc_templates(g, seed=1)
    seqs = ts.seqs()
    e = Engine(0)
    e.config(nthreads=os.cpu_count() or 8, **CFG)
    ref = None
    kinds = [('plain', 'plain', None), ('gz', 'gz', None), ('gz-1thread', 'gz', '1')] + ([('gzm', 'gzm', None), ('bgzf', 'bgzf', None)] if gzm else [])
    kinds += [('gz-%sthreads' % n, 'gz', n) for n in os.environ.get('KVQ_C3_SWEEP', '').split(',') if n]      # inflate thread sweep
    for name, key, threads in kinds:
        files = [m[key] for m in mates]
        if threads:
            os.environ['KVQ_GZ_THREADS'] = threads
        else:
            os.environ.pop('KVQ_GZ_THREADS', None)
        reps = 1 if threads == '1' else 3
        if threads != '1':
            e.findseqs(files, seqs, nplus=ts.nplus, convert=False)             # warm-up (page cache, tables)
        t = []
        for _ in range(reps):
            t0 = time.perf_counter()
            r = e.findseqs(files, seqs, nplus=ts.nplus)
            t.append(time.perf_counter() - t0)
        sig = ([tuple(h) for h in r['hits']], r['hitseqs'], r['stats']['readlengths'], r['stats']['nseqhits'],
               r['stats']['nseqbasehits'], r['stats']['records_parsed'], r['stats']['parsed'], r.coverage.tolist())
        if ref is None:
            ref = sig
        same = sig == ref
        if key != 'plain' and name != 'gz':
            pass
        dt = min(t)
        print('CONFIG3 %-10s e2e %.3f s  %.2f Gbp/s  inflated %.2f GB/s  hits %d records %d kernel_ms %.2f  == plain scan: %s' %
              (name, dt, nbases / dt / 1e9, raw_total / dt / 1e9, len(r['hits']), r['stats']['records_parsed'],
               r.timing['kernel_ms'], same), flush=True)
        assert same, name
    os.environ.pop('KVQ_GZ_THREADS', None)
    # the first npre pairs against the CPU checker, stats (incl. the float size estimate of gz input) and all
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    from util import assert_same_result
    from oracle import checker
    checker.build()
    pre = [m['pre'] for m in mates]
    got = e.findseqs(pre, seqs, nplus=ts.nplus)
    t0 = time.time()
    want = checker.findseqs(pre, seqs, nthreads=os.cpu_count() or 8, impl='oracle', **CFG)
    assert_same_result(got, want, what='config 3 prefix')
    print('CONFIG3 prefix of %d pairs (gz) == CPU checker (oracle, %d threads, %.1f s): hits %d, records %d, stats incl. total %d' %
          (npre, os.cpu_count() or 8, time.time() - t0, len(want['hits']), want['stats']['records_parsed'], want['stats']['total']), flush=True)
    for m in mates:
        for k in ('plain', 'gz', 'gzm', 'bgzf', 'pre'):
            if k in m:
                os.remove(m[k])
    os.rmdir(tmp)


if __name__ == '__main__':
    main()
