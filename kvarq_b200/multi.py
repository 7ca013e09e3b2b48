"""Multi-GPU driver of the scan: one process per GPU, every rank scans whole files
(or isolates) of the job, no data-path collective.  What is combined at the end is small and
additive -- per-template coverage, alternate-base counts, hit counters, the read-length histogram
-- with ONE ncclAllReduce inside libkvarq_b200.so (Engine.allreduce / Engine.job_allreduce).  Hit
lists stay per rank (variable length) and are merged on the host into the order the reference
gives for the unsharded file list.

torch.distributed is only the launcher-side plumbing here: it hands the NCCL unique id of rank 0
to the other ranks (any backend; the CPU tests use gloo) and provides the barrier / max used for
timing.  `host_reduce` is the CPU mirror of the library's reduce, for the gloo tests.
"""
import bisect
import os

import numpy as np


def shard(items, rank, world):
    """Round-robin deal of independent units (files, isolates) to ranks."""
    return [x for i, x in enumerate(items) if i % world == rank]


def shard_indices(n, rank, world):
    return [i for i in range(n) if i % world == rank]


def init_comm(engine, dist=None, rank=None, world=None):
    """Create the engine's NCCL communicator from the torchrun environment.  Rank 0 makes the NCCL
    unique id, `dist` (an initialised torch.distributed, any backend) broadcasts it."""
    rank = int(os.environ.get('RANK', '0')) if rank is None else rank
    world = int(os.environ.get('WORLD_SIZE', '1')) if world is None else world
    if world <= 1:
        return rank, world
    box = [engine.unique_id() if rank == 0 else None]
    if dist is None:
        raise RuntimeError('init_comm needs an initialised torch.distributed to hand out the NCCL id')
    dist.broadcast_object_list(box, src=0)
    engine.comm_init(box[0], rank, world)
    return rank, world


# ------------------------------------------------------------------ hit lists of sharded scans --
def merge_hits(rank_files, rank_results, file_sizes):
    """Per-rank results of a file list dealt to ranks -> (hits, hitseqs) of the unsharded scan
    `engine.findseqs(all files)` in the reference's nthreads=1 order.

    rank_files[r]   : indices (into the job's file list) of the files rank r scanned, in scan order
    rank_results[r] : that scan's {'hits', 'hitseqs'}; its file_pos counts from the start of the
                      rank's own concatenated stream (csrc/workhorse.c:45: fpos is never reset)
    file_sizes[i]   : inflated size of file i (gives the global file_pos base of every file)
    """
    gbase = np.concatenate([[0], np.cumsum(np.asarray(file_sizes, dtype=np.int64))])
    keyed = []
    for files, res in zip(rank_files, rank_results):
        lbase = [0]
        for f in files:
            lbase.append(lbase[-1] + int(file_sizes[f]))
        for order, (h, hs) in enumerate(zip(res['hits'], res['hitseqs'])):
            k = bisect.bisect_right(lbase, h[1]) - 1         # which of the rank's files holds the read
            f = files[k]
            g = type(h)._make((h[0], int(h[1] - lbase[k] + gbase[f]), h[2], h[3], h[4])) if hasattr(h, '_make') \
                else (h[0], int(h[1] - lbase[k] + gbase[f]), h[2], h[3], h[4])
            keyed.append((f, order, g, hs))
    keyed.sort(key=lambda x: (x[0], x[1]))                   # file order; inside a file the rank's own order
    return [k[2] for k in keyed], [k[3] for k in keyed]


# ------------------------------------------------------------- CPU mirror of the counter reduce --
def pack_counters(coverage, mutcount, nseqhits, nseqbasehits, readlengths, records):
    """One flat int64 vector per rank; layout returned for unpack_counters."""
    parts = [np.asarray(coverage, np.int64).ravel(), np.asarray(mutcount, np.int64).ravel(),
             np.asarray(nseqhits, np.int64), np.asarray(nseqbasehits, np.int64),
             np.asarray(list(readlengths) + [0] * (1024 - len(readlengths)), np.int64),
             np.asarray([records], np.int64)]
    return np.concatenate(parts), [len(p) for p in parts]


def unpack_counters(flat, layout):
    out, at = [], 0
    for n in layout:
        out.append(flat[at:at + n])
        at += n
    cov, mut, sh, sb, rls, rec = out
    longest = int(np.flatnonzero(rls).max()) if rls.any() else -1
    return {'coverage': cov, 'mutcount': mut.reshape(-1, 6), 'nseqhits': sh, 'nseqbasehits': sb,
            'readlengths': rls[:longest + 1], 'records_parsed': int(rec[0])}


def host_reduce(flat, dist):
    """Sum the packed counters over all ranks on the host (gloo): what Engine.allreduce does on the
    device with NCCL."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(flat))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.numpy()


def split_job_row(row, nseqs, cov_total):
    """One row of Engine.job_allreduce() -> dict (layout: include/kvarq_b200.h, kvq_job_begin)."""
    ns, nc = max(1, nseqs), max(1, cov_total)
    at = 2 * ns + 7 * nc
    rls = row[at:at + 1024]
    longest = int(row[at + 1027]) - 1
    return {'nseqhits': row[:nseqs], 'nseqbasehits': row[ns:ns + nseqs], 'coverage': row[2 * ns:2 * ns + cov_total],
            'mutcount': row[2 * ns + nc:2 * ns + nc + 6 * cov_total].reshape(-1, 6),
            'readlengths': rls[:min(longest + 1, 1024)], 'records_parsed': int(row[at + 1024]),
            'bases': int(row[at + 1025]), 'parsed': int(row[at + 1026])}
