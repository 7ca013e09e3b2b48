"""kvq_order_hits_static: the entry point that allocates nothing (its scratch is a file-scope device
array, no counterpart in the reference).  Collected
last on purpose.  Its output -- hits in the reference's nthreads=1 order -- is compared with the
order the existing entry point (kvq_findseqs_mem) returns, bit-exact, on two calls in a row that share
the static scratch.  The device buffers of the test come from the CUDA runtime through ctypes (the
test may allocate; the entry point may not)."""
import ctypes as C
import os

import numpy as np
import pytest

from kvarq_b200 import synth

pytestmark = pytest.mark.gpu

HIT = np.dtype([('seq_nr', '<i4'), ('seq_pos', '<i4'), ('file_pos', '<i8'), ('length', '<i4'), ('readlength', '<i4')])
COMP = bytes.maketrans(b'ACGT', b'TGCA')


def _scan():
    from kvarq_b200.engine import Engine
    g = synth.genome(20000, seed=91)
    plus = [bytes(g[p:p + 200]) for p in (1000, 5000, 9000, 13000, 17000)] + [bytes(g[p:p + 51]) for p in (3000, 7000, 11000)]
    seqs = [s.decode() for s in plus] + [s.translate(COMP)[::-1].decode() for s in plus]
    fq = np.concatenate([synth.fastq_reads(g, 30000, readlen=60, seed=92),
                         synth.fastq_reads(g, 4000, readlen=150, seed=93, first_index=50000)])
    e = Engine(0)
    e.config(maxerrors=1, minoverlap=25, minreadlength=25, Amin='.')
    r = e.findseqs_mem([fq], seqs)
    e.close()
    return r, seqs


class _Cuda(object):
    """cudaMalloc / cudaMemcpy / cudaFree of libcudart (same primary context as the library's static runtime)"""

    def __init__(self):
        self.rt = None
        for name in ('libcudart.so.12', '/usr/local/cuda/lib64/libcudart.so.12', 'libcudart.so', '/usr/local/cuda/lib64/libcudart.so'):
            try:
                self.rt = C.CDLL(name)
                break
            except OSError:
                continue
        if self.rt is None:
            pytest.skip('libcudart not found')
        self.rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        self.rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        self.rt.cudaMemset.argtypes = [C.c_void_p, C.c_int, C.c_size_t]
        self.rt.cudaFree.argtypes = [C.c_void_p]
        assert self.rt.cudaSetDevice(0) == 0
        self.bufs = []

    def alloc(self, nbytes, fill=0xEE):
        p = C.c_void_p()
        assert self.rt.cudaMalloc(C.byref(p), max(256, nbytes)) == 0
        assert self.rt.cudaMemset(p, fill, max(256, nbytes)) == 0
        self.bufs.append(p)
        return p

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.alloc(arr.nbytes)
        assert self.rt.cudaMemcpy(p, arr.ctypes.data, arr.nbytes, 1) == 0
        return p

    def download(self, p, nbytes):
        out = np.empty(nbytes, dtype=np.uint8)
        assert self.rt.cudaDeviceSynchronize() == 0
        assert self.rt.cudaMemcpy(out.ctypes.data, p, nbytes, 2) == 0
        return out

    def free(self):
        for p in self.bufs:
            self.rt.cudaFree(p)
        self.bufs = []


def _order(lib, cu, hits, seqlens, perm):
    """hits[perm] -> device -> kvq_order_hits_static -> (sorted hits, order) back on the host"""
    shuf = np.ascontiguousarray(hits[perm])
    d_in, d_len = cu.upload(shuf.view(np.uint8)), cu.upload(seqlens)
    d_out, d_ord = cu.alloc(shuf.nbytes), cu.alloc(4 * len(shuf), fill=0xFF)
    rc = lib.kvq_order_hits_static(0, d_in, len(shuf), d_len, len(seqlens), d_out, d_ord, None)
    assert rc == 0, lib.kvq_static_last_error()
    out = cu.download(d_out, shuf.nbytes).view(HIT)
    order = cu.download(d_ord, 4 * len(shuf)).view(np.uint32)
    cu.free()
    return out, order


def test_order_hits_static_matches_engine_order():
    from kvarq_b200 import _lib
    lib = _lib.load()
    cu = _Cuda()
    r, seqs = _scan()
    hits = np.array([tuple(h[k] for k in (0, 2, 1, 3, 4)) for h in r['hits']], dtype=HIT)      # Hit: seq_nr file_pos seq_pos ..
    hitseqs = np.array(r['hitseqs'], dtype=object)
    n = len(hits)
    assert 2000 < n <= 65536
    tup = [tuple(h) for h in r['hits']]
    assert len(tup) - len(set(tup)) > 10, 'no flush-right duplicates (rules b + d) in the test data'
    seqlens = np.array([len(s) for s in seqs], dtype=np.int32)
    rng = np.random.default_rng(5)
    # call 1: all hits, shuffled
    perm = rng.permutation(n)
    out, order = _order(lib, cu, hits, seqlens, perm)
    assert out.tobytes() == hits.tobytes(), 'first call: order differs from the engine\'s'
    assert list(hitseqs[perm][order]) == list(hitseqs)
    # call 2, right behind it on the same (stale) scratch: a sub-list with another shuffle
    keep = np.flatnonzero(hits['file_pos'] < np.median(hits['file_pos']))
    sub, sub_seqs = hits[keep], hitseqs[keep]
    perm2 = rng.permutation(len(sub))
    out2, order2 = _order(lib, cu, sub, seqlens, perm2)
    assert out2.tobytes() == sub.tobytes(), 'second call: order differs from the engine\'s'
    assert list(sub_seqs[perm2][order2]) == list(sub_seqs)
    # sizes beyond the static scratch are refused with a clear message, nothing is launched
    rc = lib.kvq_order_hits_static(0, None, 65537, None, 0, None, None, None)
    assert rc == _lib.KVQ_ERR_TYPE and b'KVQ_ORDER_STATIC_MAX' in lib.kvq_static_last_error()
    assert lib.kvq_order_hits_static(0, None, 0, None, 0, None, None, None) == 0
