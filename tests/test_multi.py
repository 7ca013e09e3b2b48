"""World-size-2 gloo test of the multi-GPU reduction logic on CPU: the per-rank
counters are derived with the CPU checker from two shards of one FastQ, packed, all-reduced and
must equal the counters of the unsharded scan."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _coverage(ts, hits, hitseqs):
    """Coverage.apply_hit (kvarq/analyse.py:57-78) restated for the checker's hits."""
    seqs = ts.seqs()
    lens = [len(s) for s in seqs[:ts.nplus]]
    off = np.concatenate([[0], np.cumsum(lens)])
    cov = np.zeros(off[-1], np.int64)
    for h, hs in zip(hits, hitseqs):
        seq_nr, _, seq_pos, length, _ = h
        plus = seq_nr < ts.nplus
        tpl = seq_nr if plus else seq_nr - ts.nplus
        start = max(0, seq_pos)
        for j in range(start, start + length):
            cov[off[tpl] + (j if plus else lens[tpl] - 1 - j)] += 1
    return cov


def _worker(rank, world, port, fq_shards, seqs_blob, q):
    import torch.distributed as dist
    from kvarq_b200 import multi, synth
    from oracle import checker
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    ts = seqs_blob
    cfg = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
    r = checker.findseqs(fq_shards[rank], ts.seqs(), nthreads=1, impl='oracle', **cfg)
    cov = _coverage(ts, r['hits'], r['hitseqs'])
    flat, layout = multi.pack_counters(cov, np.zeros((len(cov), 6), np.int64), r['stats']['nseqhits'],
                                       r['stats']['nseqbasehits'], r['stats']['readlengths'],
                                       r['stats']['records_parsed'])
    total = multi.host_reduce(flat, dist)
    if rank == 0:
        q.put((total, layout, [tuple(h) for h in r['hits']]))
    else:
        q.put((None, None, [tuple(h) for h in r['hits']]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_counter_reduce(tmp_path):
    import torch.multiprocessing as mp
    from kvarq_b200 import multi, synth
    from oracle import checker
    if not checker.available('oracle'):
        checker.build()
    g = synth.genome(200000, seed=21)
    ts = synth.mtbc_templates(g, seed=22)
    iso = synth.isolate(g, ts, seed=23)
    shards = []
    for i in range(2):
        fn = str(tmp_path / ('shard%d.fastq' % i))
        open(fn, 'wb').write(synth.fastq_reads(iso, 3000, readlen=150, seed=30 + i).tobytes())
        shards.append(fn)
    assert multi.shard(shards, 0, 2) == [shards[0]] and multi.shard(shards, 1, 2) == [shards[1]]
    cfg = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
    whole = [checker.findseqs(fn, ts.seqs(), nthreads=1, impl='oracle', **cfg) for fn in shards]

    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, shards, ts, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    total, layout = next((t, l) for t, l, _ in got if t is not None)
    red = multi.unpack_counters(total, layout)
    assert red['records_parsed'] == sum(w['stats']['records_parsed'] for w in whole) == 6000
    want_sh = np.sum([w['stats']['nseqhits'] for w in whole], axis=0)
    want_sb = np.sum([w['stats']['nseqbasehits'] for w in whole], axis=0)
    assert (red['nseqhits'] == want_sh).all() and (red['nseqbasehits'] == want_sb).all()
    assert int(red['coverage'].sum()) == int(want_sb.sum())          # every hit base covers one template base
    nhits = sum(len(h) for _, _, h in got)
    assert nhits == sum(len(w['hits']) for w in whole)
    rl = np.zeros(1024, np.int64)
    for w in whole:
        rl[:len(w['stats']['readlengths'])] += np.asarray(w['stats']['readlengths'], np.int64)
    assert (red['readlengths'] == rl[:len(red['readlengths'])]).all()


def test_merge_hits_equals_unsharded_scan(tmp_path):
    """Files dealt to ranks, every rank scanning its own files as one stream: the merged hit list
    (global file_pos rebuilt from the inflated file sizes, file order) must be the hit list of the
    unsharded `findseqs(all files)` in the reference's order.  ADVICE r1, multi.py:44."""
    from kvarq_b200 import multi, synth
    from oracle import checker
    if not checker.available('oracle'):
        checker.build()
    g = synth.genome(150000, seed=31)
    ts = synth.mtbc_templates(g, seed=32)
    iso = synth.isolate(g, ts, seed=33)
    files, sizes = [], []
    for i, n in enumerate((1500, 700, 2200, 900, 1300)):
        fn = str(tmp_path / ('lane%d.fastq' % i))
        data = synth.fastq_reads(iso, n, readlen=150, seed=40 + i, first_index=10000 * i).tobytes()
        open(fn, 'wb').write(data)
        files.append(fn); sizes.append(len(data))
    cfg = dict(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
    impl = 'ref' if checker.available('ref') else 'oracle'
    whole = checker.findseqs(files, ts.seqs(), nthreads=1, impl=impl, **cfg)
    assert len(whole['hits']) > 30
    for world in (2, 3):
        rank_files = [multi.shard_indices(len(files), r, world) for r in range(world)]
        results = [checker.findseqs([files[i] for i in idx], ts.seqs(), nthreads=1, impl=impl, **cfg) for idx in rank_files]
        hits, hitseqs = multi.merge_hits(rank_files, results, sizes)
        assert [tuple(h) for h in hits] == [tuple(h) for h in whole['hits']]
        assert list(hitseqs) == list(whole['hitseqs'])


@pytest.mark.gpu
def test_engine_job_rows_and_allreduce_single_rank(tmp_path):
    """The library's reduce entry points on one GPU (world 1: the collective is the identity): the job
    matrix row of a scan holds that scan's counters; with a 1-rank NCCL communicator kvq_allreduce runs
    the real ncclAllReduce and must leave the counters unchanged."""
    from kvarq_b200 import multi, synth
    from kvarq_b200.engine import Engine
    g = synth.genome(150000, seed=31)
    ts = synth.mtbc_templates(g, seed=32)
    e = Engine(0)
    e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
    rows = []
    for k in range(3):
        fq = synth.fastq_reads(synth.isolate(g, ts, seed=50 + k), 4000, readlen=150, seed=60 + k)
        r = e.findseqs_mem([fq], ts.seqs(), nplus=ts.nplus)
        if k == 0:
            e.job_begin(4)
        e.job_store(k + 1)
        rows.append(r)
    m, _ = e.job_allreduce()
    cov_total = len(rows[0].coverage)
    assert not m[0].any()
    for k, r in enumerate(rows):
        d = multi.split_job_row(m[k + 1], 2 * ts.nplus, cov_total)
        assert d['coverage'].tolist() == r.coverage.tolist() and d['mutcount'].tolist() == r.mutcount.tolist()
        assert tuple(d['nseqhits']) == r['stats']['nseqhits'] and tuple(d['nseqbasehits']) == r['stats']['nseqbasehits']
        assert tuple(d['readlengths']) == r['stats']['readlengths'] and d['records_parsed'] == 4000
    e.comm_init(Engine.unique_id(), 0, 1)
    assert e.comm_info()['world'] == 1 and e.comm_info()['nccl_version'] > 20000
    r = e.findseqs_mem([fq], ts.seqs(), nplus=ts.nplus)
    red = e.allreduce()
    assert red.coverage.tolist() == r.coverage.tolist() and red['stats'] == r['stats']
    assert red.timing['reduce_ms'] > 0
    # the one-call form (kvq_scan_staged_reduce: the collective enqueued behind the scan kernels, tail of the row
    # written on the device) gives the same hits and the same counters -- also when the batch has to be run
    # again because the survivor queue overflowed (the collective is repeated then)
    st = e.stage([fq])
    plain = e.scan_staged(st, ts.seqs(), nplus=ts.nplus)
    fused = e.scan_staged(st, ts.seqs(), nplus=ts.nplus, reduce=True)
    assert fused['hits'] == plain['hits'] == r['hits'] and fused['hitseqs'] == plain['hitseqs']
    assert fused['stats'] == plain['stats'] and fused.coverage.tolist() == plain.coverage.tolist()
    assert fused.mutcount.tolist() == plain.mutcount.tolist() and fused.timing['reduce_ms'] > 0
    e.close()


@pytest.mark.gpu
def test_fused_reduce_repeats_the_collective_after_a_rerun(monkeypatch):
    """KVQ_SURV_CAP=16: the first pass overflows the survivor queue, the rank runs the batch again; its row in the
    first sum carries the 'incomplete' flag, so the collective is made a second time with the final counters"""
    from kvarq_b200 import synth
    from kvarq_b200.engine import Engine
    monkeypatch.setenv('KVQ_SURV_CAP', '16')
    g = synth.genome(150000, seed=31)
    ts = synth.mtbc_templates(g, seed=32)
    fq = synth.fastq_reads(synth.isolate(g, ts, seed=52), 4000, readlen=150, seed=62)
    e = Engine(0)
    e.config(maxerrors=2, minoverlap=25, minreadlength=25, Amin='.')
    e.comm_init(Engine.unique_id(), 0, 1)
    st = e.stage([fq])
    fused = e.scan_staged(st, ts.seqs(), nplus=ts.nplus, reduce=True)
    plain = e.scan_staged(st, ts.seqs(), nplus=ts.nplus)
    assert len(fused['hits']) > 50 and fused['hits'] == plain['hits'] and fused['stats'] == plain['stats']
    assert fused.coverage.tolist() == plain.coverage.tolist() and fused.timing['launches'] >= plain.timing['launches']
    e.close()
