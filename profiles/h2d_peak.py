#!/usr/bin/env python3
"""Host->device copy ceiling of the box (development tool): N processes, one per GPU,
each copying a 0.98 GB page-locked buffer to its GPU again and again (plain cudaMemcpyAsync through
torch, 64 MiB pieces like the engine's batches); all start together, the aggregate GB/s is what `e2e` at
N GPUs can reach at best.  Also prints the NUMA node of every GPU and the CPU count.
usage: h2d_peak.py [N ...]     (default: 1, then 2, 4, 8 as far as the box has GPUs)"""
import os
import sys
import time

import torch
import torch.multiprocessing as mp

NBYTES = 979 * 1000 * 1000
PIECE = 64 << 20


def worker(rank, world, start_at, q):
    torch.cuda.set_device(rank)
    a = torch.empty(NBYTES, dtype=torch.uint8).pin_memory()
    b = torch.empty(NBYTES, dtype=torch.uint8, device='cuda')
    for _ in range(2):
        b.copy_(a, non_blocking=True)
    torch.cuda.synchronize()
    while time.time() < start_at:
        pass
    reps = 10
    t = time.perf_counter()
    for _ in range(reps):
        for o in range(0, NBYTES, PIECE):
            b[o:o + PIECE].copy_(a[o:o + PIECE], non_blocking=True)
    torch.cuda.synchronize()
    q.put((rank, reps * NBYTES / (time.perf_counter() - t) / 1e9))


def main():
    ngpu = torch.cuda.device_count()
    sizes = [int(x) for x in sys.argv[1:]] or [n for n in (1, 2, 4, 8) if n <= ngpu]
    print('cpus %d, gpus %d' % (os.cpu_count(), ngpu))
    for i in range(ngpu):
        try:
            bus = torch.cuda.get_device_properties(i).pci_bus_id
            node = open('/sys/bus/pci/devices/0000:%02x:00.0/numa_node' % bus).read().strip()
        except Exception:
            node = '?'
        print('gpu %d numa node %s' % (i, node))
    ctx = mp.get_context('spawn')
    for n in sizes:
        q = ctx.Queue()
        start_at = time.time() + 10 + n
        ps = [ctx.Process(target=worker, args=(r, n, start_at, q)) for r in range(n)]
        for p in ps:
            p.start()
        res = sorted(q.get() for _ in ps)
        for p in ps:
            p.join()
        print('H2D_PEAK n=%d aggregate %.1f GB/s per gpu %s' % (n, sum(r[1] for r in res), ' '.join('%.1f' % r[1] for r in res)), flush=True)


if __name__ == '__main__':
    main()
