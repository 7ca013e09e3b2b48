#!/usr/bin/env python3
"""Host-side ceiling of the file path (development tool): N threads pread() a page-cache file into one
buffer (plain and CUDA page-locked), no GPU work.  usage: read_peak.py [MB]"""
import os, sys, time, ctypes as C
from concurrent.futures import ThreadPoolExecutor
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
fn = '/dev/shm/kvq_read_peak.bin'
a = np.random.default_rng(1).integers(0, 255, mb << 20, dtype=np.uint8)
open(fn, 'wb').write(a)
fd = os.open(fn, os.O_RDONLY)
part = 2 << 20

def sweep(buf, tag):
    mv = memoryview(buf)
    for th in (1, 2, 4, 8, 16, 32):
        def job(i):
            os.preadv(fd, [mv[i * part:(i + 1) * part]], i * part)
        best = 1e9
        with ThreadPoolExecutor(th) as ex:
            for _ in range(3):
                t = time.perf_counter(); list(ex.map(job, range(len(buf) // part))); best = min(best, time.perf_counter() - t)
        print('%s threads %2d: %.1f GB/s' % (tag, th, len(buf) / best / 1e9), flush=True)

sweep(np.empty(mb << 20, dtype=np.uint8), 'pageable')
try:
    from kvarq_b200 import _lib
    lib = _lib.load()
    p = lib.kvq_host_alloc(mb << 20)
    pin = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(mb << 20,))
    sweep(pin, 'pinned  ')
except Exception as e:
    print('no pinned test:', e)
os.remove(fn)
