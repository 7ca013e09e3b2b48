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
    total = multi.all_reduce_counters(flat, dist)
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
