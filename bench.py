#!/usr/bin/env python
"""bench.py -- Gbp scanned/sec of the read-scanning hot path on N B200s (BASELINE.json metric).

Workload (BASELINE.json configs[1]): synthetic H37Rv-like 4.4 Mb genome, 100x coverage, 3 M
single-end 150 bp Illumina-profile reads (~0.98 GB FastQ) against the MTBC-shaped template set
(132 templates -> 264 seqs incl. reverse complements), product defaults maxerrors=2
minoverlap=25 minreadlength=25 Q13 (Amin='.').  One "step" = one full pass of the scan over
that FastQ.  With N GPUs every rank scans its own isolate of that shape (weak scaling, no
data-path collective); per-template coverage / hit counters are summed with one NCCL
all-reduce at the end.

  value     : bases scanned / time, FastQ resident in HBM (kvq_scan_staged), hits + fused
              coverage produced and read back every step
  e2e       : same scan through kvq_findseqs_mem on HOST buffers: host -> pinned -> HBM copies
              and the device -> host result copies inside the timed region
  roofline  : algorithmic bytes (= FastQ bytes, each read once) / CUDA-event time of the scan
              kernels, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline / --impl reference : the reference's own C engine (oracle/_ref, csrc/workhorse.c
              compiled unmodified) on the host cores, bounded sample of the same workload
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
READS = int(os.environ.get('KVQ_BENCH_READS', 3000000))
READLEN = 150


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# stdout carries the one JSON line and nothing else: libraries that write to fd 1 on their own (NCCL's
# version banner) are sent to stderr, the line goes out through a private copy of the original stdout
_OUT = None


def claim_stdout():
    global _OUT
    if _OUT is None:
        sys.stdout.flush()
        _OUT = os.fdopen(os.dup(1), 'w')
        os.dup2(2, 1)


def emit(obj):
    out = _OUT or sys.stdout
    out.write(json.dumps(obj) + '\n')
    out.flush()


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        return float(p['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


class ClockSampler(object):
    """nvidia-smi clocks + throttle reasons during the timed regions (B200_PROFILING.md).  A separate
    process (`nvidia-smi -lms`) so that sampling never competes with the timed loop for the GIL;
    only samples stamped inside [mark_begin, mark_end] windows are used."""
    Q = ('timestamp,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.windows = []
        self.t0 = None
        self.proc = None
        if index is None:                   # ranks > 0: one sampler per box (rank 0's GPU) is enough, and
            return                          # concurrent nvidia-smi pollers slow every rank's CUDA calls
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '20'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def start(self):
        pass

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.windows.append((self.t0, time.time()))

    def summary(self):
        import datetime
        rows = []
        if self.proc is not None:
            time.sleep(0.05)
            self.proc.terminate()
            try:
                out = self.proc.communicate(timeout=5)[0]
            except Exception:
                out = ''
            for line in out.strip().split('\n'):
                f = [x.strip() for x in line.split(',')]
                if len(f) < 7:
                    continue
                try:
                    ts = datetime.datetime.strptime(f[0], '%Y/%m/%d %H:%M:%S.%f').timestamp()
                except Exception:
                    continue
                rows.append((ts, f[1:]))
        inside = [f for ts, f in rows if any(a - 0.02 <= ts <= b + 0.02 for a, b in self.windows)]
        use = inside or [f for _, f in rows]
        sm = sorted(int(f[0]) for f in use if f[0].isdigit())
        mx = [int(f[1]) for f in use if f[1].isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = sorted({n for f in use for n, v in zip(names, f[2:6]) if v == 'Active'})
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': reasons, 'samples': len(use), 'samples_inside_timed_regions': len(inside)}


def make_workload(rank):
    from kvarq_b200 import synth
    t0 = time.time()
    g = synth.genome()
    ts = synth.mtbc_templates(g, seed=1)
    iso = synth.isolate(g, ts, seed=7 + rank)
    fq = synth.fastq_reads(iso, READS, readlen=READLEN, seed=2 + 1000 * rank)
    log('[rank %d] workload: %d reads, %.1f MB FastQ, %d seqs, %.1f s' %
        (rank, READS, fq.nbytes / 1e6, 2 * ts.nplus, time.time() - t0))
    return ts, fq


def cpu_scan(fq_bytes, seqs, nthreads, tmpdir, tag):
    """Times the reference CPU engine (oracle/_ref, else the port) on a FastQ sample."""
    from oracle import checker
    if not checker.available('oracle'):
        checker.build()
    impl = 'ref' if checker.available('ref') else 'oracle'
    fn = os.path.join(tmpdir, 'kvq_cpu_%s_%d.fastq' % (tag, os.getpid()))
    with open(fn, 'wb') as f:
        f.write(fq_bytes)
    try:
        t0 = time.time()
        r = checker.findseqs(fn, seqs, nthreads=nthreads, impl=impl, want_hits=False, **CFG)
        dt = time.time() - t0
    finally:
        os.remove(fn)
    return dt, r, impl


def prefix_parity(eng, pre, seqs, nplus, cpu_result, impl, cores, tmpdir):
    """The CPU leg's per-template counters / histogram / record count against a GPU scan of the
    same prefix (outside every timed region).  The reference's own C races with nthreads > 1
    (csrc/workhorse.c:953-955: a chunk-boundary record can be scanned twice), so when that run
    differs the deterministic port (oracle/kvarq_oracle.c, pinned to the reference at nthreads=1)
    is run on the same bytes and decides."""
    keys = ('nseqhits', 'nseqbasehits', 'readlengths', 'records_parsed')
    g = eng.findseqs_mem([pre], seqs, nplus=nplus, want_hits=False)['stats']
    same = lambda st: all(list(g[k]) == list(st[k]) if k != 'records_parsed' else g[k] == st[k] for k in keys)
    if same(cpu_result['stats']):
        return 'equal'
    if impl == 'ref':
        from oracle import checker
        fn = os.path.join(tmpdir, 'kvq_cpu_parity_%d.fastq' % os.getpid())
        with open(fn, 'wb') as f:
            f.write(pre.tobytes())
        try:
            o = checker.findseqs(fn, seqs, nthreads=cores, impl='oracle', want_hits=False, **CFG)
        finally:
            os.remove(fn)
        if same(o['stats']):
            return ('equal (vs the deterministic port; the reference C at nthreads=%d raced: records_parsed %d vs %d)'
                    % (cores, cpu_result['stats']['records_parsed'], g['records_parsed']))
    return 'DIFFERS'


def sample_bytes(fq, nreads):
    """First nreads records of the (fixed-width per decade) synthetic FastQ."""
    import numpy as np
    nl = np.flatnonzero(fq[:min(fq.nbytes, nreads * 400 + 4096)] == 10)
    return fq[:int(nl[4 * nreads - 1]) + 1]


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's own C engine on the host cores, rank 0 only."""
    if rank != 0:
        return
    from kvarq_b200 import synth
    import numpy as np
    cores = os.cpu_count() or 1
    ts, fq = None, None
    g = synth.genome()
    ts = synth.mtbc_templates(g, seed=1)
    # ~0.75 ms per read and core (BASELINE.md probe): size one step to ~8 s of wall clock
    per_step = int(min(READS, max(2000, 8.0 * cores / 0.00075 * 0.6)))
    total = per_step * (args.steps + args.warmup)
    fq = synth.fastq_reads(synth.isolate(g, ts, seed=7), total, readlen=READLEN, seed=2)
    tmp = '/dev/shm' if os.path.isdir('/dev/shm') else '/tmp'
    seqs = ts.seqs()
    nl = np.flatnonzero(fq == 10)
    cut = lambda i: 0 if i == 0 else int(nl[4 * per_step * i - 1]) + 1
    times, impl = [], None
    for i in range(args.warmup + args.steps):
        dt, r, impl = cpu_scan(fq[cut(i):cut(i + 1)].tobytes(), seqs, cores, tmp, 'ref')
        if i >= args.warmup:
            times.append(dt)
        log('[reference] step %d: %d reads in %.2f s' % (i, per_step, dt))
    t = sum(times)
    bases = per_step * READLEN * args.steps
    val = bases / t / 1e9
    kind = 'reference' if impl == 'ref' else 'port'
    sample = '%d steps x first-%d-record slices of the 3M-read workload, nthreads=%d' % (args.steps, per_step, cores)
    emit({
        'impl': 'reference', 'metric': 'Gbp scanned/sec', 'value': val, 'unit': 'Gbp/s', 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * t / args.steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic',
        'config': workload_config(),
        'cpu_baseline': {'value': val, 'unit': 'Gbp/s', 'cores': cores, 'kind': kind, 'sample': sample},
        'e2e': {'value': val, 'unit': 'Gbp/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    })


def run_config5(args, rank, world, local, dist, barrier, allmax, allsum):
    """BASELINE.json configs[4]: a batch of synthetic isolates (seeds 500.., the config-2 recipe with a
    random genotype each) dealt round-robin to the ranks.  Every isolate goes through the
    end-to-end call on host buffers (page-locked, H2D inside the timed region); its counters are
    stored as one row of an [isolates][row] int64 matrix in HBM; ONE ncclAllReduce over the matrix at
    the end gives every rank the coverage / hit counters of every isolate.  Timed: the scans + row
    stores of a rank (max over ranks) + the reduce; the reads are generated on the host between the
    scans, outside the timed regions."""
    import numpy as np
    from kvarq_b200 import multi, synth
    from kvarq_b200.engine import Engine
    n_iso, n_reads = args.isolates, args.reads_per_isolate
    g = synth.genome()
    ts = synth.mtbc_templates(g, seed=1)
    seqs = ts.seqs()
    eng = Engine(local)
    eng.config(nthreads=max(1, (os.cpu_count() or 8) // max(1, world)), **CFG)
    if dist is not None:
        multi.init_comm(eng, dist, rank, world)
    mine = multi.shard_indices(n_iso, rank, world)
    pin = None
    scan_s, bases, nbytes, nhits, kernel_ms = 0.0, 0, 0, 0, 0.0
    sampler = ClockSampler(local if rank == 0 else None)
    for k, i in enumerate(mine):
        t0 = time.time()
        iso = synth.isolate(g, ts, fraction=0.3, seed=500 + i)
        fq = synth.fastq_reads(iso, n_reads, readlen=READLEN, seed=500 + i)
        if pin is None:
            pin = eng.pinned_empty(int(fq.nbytes * 1.02) + 4096)
        buf = pin[:fq.nbytes]
        buf[:] = fq
        tgen = time.time() - t0
        if k == 0:
            eng.findseqs_mem([buf[:min(fq.nbytes, 32 << 20)]], seqs, nplus=ts.nplus, want_hits=False)     # warm-up: tables, buffers
            if world > 1:
                eng.allreduce(convert='timing')         # warm-up of the communicator (NCCL connects lazily)
            eng.job_begin(n_iso)
            if world > 1:
                eng.job_allreduce()                     # same for a collective of the matrix's size (all zeros still)
        sampler.mark_begin()
        t0 = time.perf_counter()
        r = eng.findseqs_mem([buf], seqs, nplus=ts.nplus, want_hits=False)
        eng.job_store(i)
        scan_s += time.perf_counter() - t0
        sampler.mark_end()
        bases += r.timing['bases']; nbytes += r.timing['stream_bytes']; nhits += r.nhits; kernel_ms += r.timing['kernel_ms']
        assert r['stats']['records_parsed'] == n_reads
        log('[rank %d] isolate %d: generated in %.1f s, scanned in %.1f ms, %d hits' % (rank, i, tgen, r.timing['wall_ms'], r.nhits))
    barrier()
    t0 = time.perf_counter()
    m, reduce_ms = eng.job_allreduce()
    barrier()
    red_s = time.perf_counter() - t0
    scan_max = allmax(scan_s)
    total_s = scan_max + allmax(red_s)
    total_bases = allsum(bases)
    h2d_total, hits_total = allsum(nbytes), allsum(nhits)
    cov_total = sum(len(x) for x in ts.plus)
    for i in range(n_iso):                       # every rank now holds every isolate's counters
        d = multi.split_job_row(m[i], len(seqs), cov_total)
        assert d['records_parsed'] == n_reads, 'isolate %d: %d records' % (i, d['records_parsed'])
        assert int(d['coverage'].sum()) == int(d['nseqbasehits'].sum())
    if rank == 0:
        emit({
            'metric': 'Gbp scanned/sec', 'value': total_bases / total_s / 1e9, 'unit': 'Gbp/s', 'n_gpus': world,
            'steps': 1, 'warmup': 0, 'ms_per_step': 1e3 * total_s, 'higher_is_better': True, 'scaling': 'strong',
            'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic',
            'config': {'workload': 'configs[4]: batch of %d synthetic isolates (seeds 500..%d, %d x %d bp reads each, %.1f Gbp) dealt to %d '
                                   'ranks, MTBC templates (264 seqs), end-to-end from page-locked host buffers, per-isolate counters '
                                   '[%d][%d] int64 reduced with one ncclAllReduce' % (n_iso, 499 + n_iso, n_reads, READLEN,
                                                                                   total_bases / 1e9, world, n_iso, m.shape[1]),
                       'isolates': n_iso, 'reads_per_isolate': n_reads},
            'e2e': {'value': total_bases / total_s / 1e9, 'unit': 'Gbp/s', 'h2d_bytes_per_step': int(h2d_total),
                    'd2h_bytes_per_step': int(m.nbytes * world)},
            'reduce': {'ms': reduce_ms, 'wall_ms': 1e3 * red_s, 'int64_counters': int(m.size), 'in_timed_region': True},
            'scan_ms_max_rank': 1e3 * scan_max, 'kernel_ms_sum_rank0': kernel_ms, 'hits_total': int(hits_total),
            'gpu_launches': 3 * len(mine) * 16, 'clocks': sampler.summary(),
        })
    eng.close()


def workload_config():
    return {'workload': 'configs[1]: synthetic H37Rv-like 4.4 Mb genome, 100x, %d x %d bp SE reads vs MTBC '
                        'templates (132 -> 264 seqs), maxerrors=2 minoverlap=25 minreadlength=25 Q13' % (READS, READLEN),
            'reads_per_gpu': READS, 'readlen': READLEN, 'nseqs': 264,
            'l2_policy': 'input (0.98 GB per pass) larger than L2; no flush needed'}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--e2e-steps', type=int, default=0, help='0 = min(steps, 3)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--config', type=int, default=2, choices=[2, 5],
                    help='2: the headline workload (default); 5: batch of isolates over all ranks + one NCCL reduce')
    ap.add_argument('--isolates', type=int, default=96)
    ap.add_argument('--reads-per-isolate', type=int, default=2080000, help='config 5: 312 Mbp per isolate at 150 bp')
    args = ap.parse_args()
    claim_stdout()
    args.warmup = max(args.warmup, 3) if args.impl == 'b200' else args.warmup

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))

    if args.impl == 'reference':
        run_reference_arm(args, rank, world)
        return

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    def barrier():
        if dist is not None:
            dist.barrier()

    def allmax(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    from kvarq_b200.engine import Engine
    if args.config == 5:
        run_config5(args, rank, world, local, dist, barrier, allmax, allsum)
        if dist is not None:
            dist.destroy_process_group()
        return
    ts, fq = make_workload(rank)
    seqs = ts.seqs()
    eng = Engine(local)
    eng.config(nthreads=max(1, (os.cpu_count() or 8) // max(1, world)), **CFG)
    if dist is not None:
        from kvarq_b200 import multi
        multi.init_comm(eng, dist, rank, world)         # NCCL communicator inside libkvarq_b200.so
        log('[rank %d] library communicator: %s' % (rank, eng.comm_info()))

    # ---- value: FastQ resident in HBM ------------------------------------------------------
    staged = eng.stage([fq])
    for _ in range(args.warmup):
        r = eng.scan_staged(staged, seqs, nplus=ts.nplus, want_hits=True, convert=False, reduce=world > 1)
    sampler = ClockSampler(local if rank == 0 else None)
    sampler.start()
    barrier()
    sampler.mark_begin()
    t0 = time.perf_counter()
    kernel_ms, launches, d2h, call_ms, reduce_ms = 0.0, 0, 0, 0.0, 0.0
    for _ in range(args.steps):
        # world > 1: the per-template counters of all ranks are summed inside the call -- one ncclAllReduce per step, enqueued
        # behind the scan kernels on its own stream (kvq_scan_staged_reduce); the call syncs once, hits and sums are on the host
        r = eng.scan_staged(staged, seqs, nplus=ts.nplus, want_hits=True, convert='timing', reduce=world > 1)
        kernel_ms += r.timing['kernel_ms']; launches += r.timing['launches']; d2h += r.timing['d2h_bytes']; call_ms += r.timing['wall_ms']
        if world > 1:
            reduce_ms += r.timing['reduce_ms']
    barrier()
    dt = allmax(time.perf_counter() - t0)
    sampler.mark_end()
    nhits_timed, nrec_timed = r.nhits, r['stats']['records_parsed']
    r = eng.scan_staged(staged, seqs, nplus=ts.nplus, want_hits=True, convert=False)   # full result for the reduce below
    assert r.nhits == nhits_timed and world * r['stats']['records_parsed'] == nrec_timed      # (the timed steps hold the sums over the ranks)
    bases = r.timing['bases']
    stream_bytes = r.timing['stream_bytes']
    total_bases = allsum(bases * args.steps)
    value = total_bases / dt / 1e9
    kms = allmax(kernel_ms / args.steps)
    log('[rank %d] resident: %.3f ms/step wall, %.3f ms/step inside the C call, %.3f ms/step kernels, %d hits, %d records, %d candidates' %
        (rank, 1e3 * dt / args.steps, call_ms / args.steps, kernel_ms / args.steps, r.nhits, r['stats']['records_parsed'], r.timing['candidates']))

    # ---- the reduced counters of one more step, checked (the timed loop above already reduced every step)
    reduce_info = None
    if world > 1:
        import numpy as np
        red = eng.allreduce()                           # the two-call form (kvq_allreduce) of the same reduce
        assert red['stats']['records_parsed'] == world * r['stats']['records_parsed'] == nrec_timed
        assert int(red.coverage.sum()) == int(np.asarray(red['stats']['nseqbasehits']).sum())
        nwords = 2 * len(seqs) + 7 * len(red.coverage) + 1029
        reduce_info = {'ms_per_step': allmax(reduce_ms / args.steps), 'int64_counters': nwords, 'in_timed_region': True,
                       'where': 'libkvarq_b200.so kvq_scan_staged_reduce: ncclAllReduce(int64, sum) of the device accumulators, enqueued behind the scan kernels on its own stream, beside the ordering of the hits'}
        log('[rank %d] NCCL all-reduce of %d int64 counters inside every step: %.3f ms/step, %d records, total coverage %d' %
            (rank, nwords, reduce_ms / args.steps, red['stats']['records_parsed'], int(red.coverage.sum())))

    # ---- e2e: the reference's own call, engine.findseqs(filename) (csrc/workhorse.c:1249), on a FastQ
    # file in the page cache (/dev/shm): file -> pinned batches -> HBM -> results, everything inside the
    # timed region.  e2e_mem: the same scan from a page-locked host buffer (kvq_findseqs_mem).
    e2e_steps = args.e2e_steps or min(args.steps, 3)
    tmp = '/dev/shm' if os.path.isdir('/dev/shm') else '/tmp'
    fq_file = os.path.join(tmp, 'kvq_bench_%d_%d.fastq' % (os.getpid(), rank))
    with open(fq_file, 'wb') as f:
        f.write(memoryview(fq))
    try:
        eng.findseqs(fq_file, seqs, nplus=ts.nplus, convert=False)                      # warm-up
        barrier()
        sampler.mark_begin()
        t0 = time.perf_counter()
        h2d = d2h_e = 0
        for _ in range(e2e_steps):
            q = eng.findseqs(fq_file, seqs, nplus=ts.nplus, convert=False)
            h2d += q.timing['h2d_bytes']; d2h_e += q.timing['d2h_bytes']
        barrier()
        dt_e = allmax(time.perf_counter() - t0)
        sampler.mark_end()
    finally:
        os.remove(fq_file)
    e2e = allsum(q.timing['bases'] * e2e_steps) / dt_e / 1e9
    assert q.nhits == r.nhits and q['stats']['records_parsed'] == r['stats']['records_parsed']
    log('[rank %d] e2e (file): %.1f ms/step' % (rank, 1e3 * dt_e / e2e_steps))
    fq_pin = eng.pinned_copy(fq)                                                    # inputs in pinned host memory
    eng.findseqs_mem([fq_pin], seqs, nplus=ts.nplus, want_hits=False)               # warm-up
    barrier()
    sampler.mark_begin()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        qm = eng.findseqs_mem([fq_pin], seqs, nplus=ts.nplus, want_hits=False)
    barrier()
    dt_m = allmax(time.perf_counter() - t0)
    sampler.mark_end()
    clocks = sampler.summary()
    e2e_mem = allsum(qm.timing['bases'] * e2e_steps) / dt_m / 1e9
    assert qm.nhits == r.nhits and qm['stats']['records_parsed'] == r['stats']['records_parsed']
    log('[rank %d] e2e (pinned buffer): %.1f ms/step' % (rank, 1e3 * dt_m / e2e_steps))

    # ---- CPU baseline on rank 0 at N=1 -------------------------------------------------------
    cpu = None
    parity_prefix = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        n = int(min(READS, max(5000, 15.0 * cores / 0.00075 * 0.6)))
        tmp = '/dev/shm' if os.path.isdir('/dev/shm') else '/tmp'
        pre = sample_bytes(fq, n)
        cdt, cr, impl = cpu_scan(pre.tobytes(), seqs, cores, tmp, 'base')
        cpu = {'value': n * READLEN / cdt / 1e9, 'unit': 'Gbp/s', 'cores': cores,
               'kind': 'reference' if impl == 'ref' else 'port',
               'sample': 'first %d of %d records, nthreads=%d, %.1f s' % (n, READS, cores, cdt)}
        log('[cpu] %s' % cpu)
        parity_prefix = prefix_parity(eng, pre, seqs, ts.nplus, cr, impl, cores, tmp)
        log('[parity] prefix of %d records: %s' % (n, parity_prefix))

    if rank == 0:
        peak, peak_src = peaks()
        achieved = stream_bytes / (kms * 1e-3) / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json')))['dram_bytes_per_launch']
        except Exception:
            pass
        emit({
            'metric': 'Gbp scanned/sec', 'value': value, 'unit': 'Gbp/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': 1e3 * dt / args.steps, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic',
            'config': workload_config(),
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                         'frac': achieved / peak, 'traffic': traffic, 'peak_source': peak_src,
                         'kernel': 'kvq_scan_kernel', 'kernel_ms_per_step': kms,
                         'algorithmic_bytes_per_launch': stream_bytes},
            'cpu_baseline': cpu, 'parity_prefix': parity_prefix,
            'e2e': {'value': e2e, 'unit': 'Gbp/s', 'h2d_bytes_per_step': h2d // e2e_steps,
                    'd2h_bytes_per_step': d2h_e // e2e_steps, 'steps': e2e_steps,
                    'path': 'engine.findseqs(file in /dev/shm): kvq_findseqs, file -> pinned batches -> HBM -> results'},
            'e2e_mem': {'value': e2e_mem, 'unit': 'Gbp/s', 'path': 'kvq_findseqs_mem on a page-locked host buffer'},
            'reduce': reduce_info,
            'gpu_launches': launches, 'clocks': clocks,
            'hits_per_step': r.nhits, 'records_per_step': r['stats']['records_parsed'],
        })
    eng.release_pinned(fq_pin)
    staged.free()
    eng.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
