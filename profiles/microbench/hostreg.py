#!/usr/bin/env python3
"""Can the file path DMA straight out of the page cache?  Times cudaHostRegister / H2D copy /
cudaHostUnregister on 64 MiB windows of an mmap'ed /dev/shm file (development tool)."""
import ctypes as C, mmap, os, time
import numpy as np
rt = C.CDLL('libcudart.so.12')
rt.cudaHostRegister.argtypes = [C.c_void_p, C.c_size_t, C.c_uint]
rt.cudaHostUnregister.argtypes = [C.c_void_p]
rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
fn = '/dev/shm/hostreg.bin'
n = 1 << 30
open(fn, 'wb').write(np.random.default_rng(1).integers(0, 255, n, dtype=np.uint8))
fd = os.open(fn, os.O_RDONLY)
d = C.c_void_p(); assert rt.cudaMalloc(C.byref(d), n) == 0
for flags, prot, name in ((0, mmap.PROT_READ | mmap.PROT_WRITE, 'private rw, flags 0'), (8, mmap.PROT_READ, 'shared ro, read-only flag')):
    m = mmap.mmap(fd, n, flags=(mmap.MAP_PRIVATE if flags == 0 else mmap.MAP_SHARED), prot=prot)
    base = C.addressof(C.c_char.from_buffer(m)) if prot & mmap.PROT_WRITE else None
    if base is None:
        a = np.frombuffer(m, dtype=np.uint8); base = a.ctypes.data
    W = 64 << 20
    treg = tcp = tun = 0.0; ok = True
    for o in range(0, n, W):
        t = time.perf_counter(); rc = rt.cudaHostRegister(base + o, W, flags); treg += time.perf_counter() - t
        if rc != 0:
            print(name, 'cudaHostRegister failed rc', rc); rt.cudaGetLastError(); ok = False; break
        t = time.perf_counter(); rt.cudaMemcpyAsync(d.value + o, base + o, W, 1, None); rt.cudaDeviceSynchronize(); tcp += time.perf_counter() - t
        t = time.perf_counter(); rt.cudaHostUnregister(base + o); tun += time.perf_counter() - t
    if ok:
        k = n // W
        print('%s: per 64 MiB window: register %.2f ms, H2D %.2f ms (%.1f GB/s), unregister %.2f ms' % (name, 1e3 * treg / k, 1e3 * tcp / k, W / (tcp / k) / 1e9, 1e3 * tun / k))
os.remove(fn)
