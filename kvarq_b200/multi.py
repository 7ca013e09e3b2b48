"""Multi-GPU plumbing of the scan: the read stream shards by file / chunk, one
process per GPU, no data-path collective.  What is combined at the end is small and additive --
per-template coverage, alternate-base counts, hit counters, the read-length histogram -- with one
all-reduce (NCCL on GPUs; gloo in the CPU tests).  Hit lists stay per rank (variable length) and
are merged on the host in stream order."""
import numpy as np


def shard(items, rank, world):
    """Round-robin deal of independent units (files, isolates, chunks) to ranks."""
    return [x for i, x in enumerate(items) if i % world == rank]


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


def all_reduce_counters(flat, dist, device=None):
    """Sum the packed counters over all ranks (torch.distributed is plumbing only)."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(flat))
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()


def merge_hits(per_rank):
    """Hit lists of the ranks -> one list in stream order (file_pos is global per stream)."""
    allh = [h for hits in per_rank for h in hits]
    return sorted(allh, key=lambda h: (h[1], h[0]))
